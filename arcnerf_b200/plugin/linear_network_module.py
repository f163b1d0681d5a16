# -*- coding: utf-8 -*-
"""The vanilla-NeRF networks: arcnerf/models/base_modules/geo_rad_model/linear_network_module.py
GeoNet (:17-197) and RadianceNet (:234-335) with the same constructor arguments, layer structure (`layers.{i}.weight /
bias`, so reference checkpoints load) and forward contracts.

Every dense layer here runs on the library's own fp32 kernels (plugin/small_ops.py -> csrc/small_gemm.cu: no torch.nn.Linear,
no cuBLAS) and is differentiable to second order, which is what the SDF models need (normals by forward_with_grad, eikonal
loss).  These general mirrors serve the narrow networks (the 64-wide, weight-normed softplus SDF net of the NeuS-NGP
configuration) and any shape of the classes; the 8 x 256 vanilla-NeRF / NeRF++ pair has its own tensor-core path,
plugin/nerf_mlp.py FusedNeRFNets (csrc/dense_chain.cu), which loads the same state dicts.  Siren layers are not mirrored."""
import math

import numpy as np
import torch
import torch.nn as nn

from . import small_ops
from .registry import ENCODER_REGISTRY, MODULE_REGISTRY, cfg_to_dict


def get_activation(cfg=None, default='relu'):
    """arcnerf/models/base_modules/activation.py: {type: ReLU | Softplus(beta) | Sigmoid | ...}; None -> default"""
    d = cfg_to_dict(cfg)
    t = (d.get('type') or default).lower()
    if t == 'relu':
        return nn.ReLU()
    if t == 'softplus':
        return _Softplus(beta=d.get('beta', 100))
    if t == 'sigmoid':
        return nn.Sigmoid()
    if t == 'leakyrelu':
        return nn.LeakyReLU(d.get('slope', 0.01))
    raise NotImplementedError('activation {}'.format(t))


def build_encoder(cfg):
    """arcnerf/models/base_modules/encoding/__init__.py:26-51 -> (embed_fn, input_ch, embed_freq); None -> identity"""
    from . import freq_encoder, hashgrid_encoder, sh_encoder  # noqa: F401  (registration)
    d = cfg_to_dict(cfg)
    if not d or d.get('type') is None:
        d = dict(type='FreqEmbedder', input_dim=d.get('input_dim', 3), n_freqs=0)
    enc = ENCODER_REGISTRY.get(d['type'])(**{k: v for k, v in d.items() if k != 'type'})
    return enc, d.get('input_dim', 3), d.get('n_freqs', 0)


class _Softplus(nn.Module):
    """nn.Softplus(beta) on the library's kernel (differentiable twice)"""

    def __init__(self, beta=100):
        super().__init__()
        self.beta = float(beta)

    def forward(self, x):
        return small_ops.softplus(x, self.beta)


class Linear(nn.Linear):
    """nn.Linear's parameters (weight (out,in), bias; so that the reference's state dicts load and nn.utils.weight_norm wraps it)
    with the product on arcb200_gemm_nt"""

    def forward(self, x):
        return small_ops.linear(x, self.weight, self.bias)


class DenseLayer(Linear):
    """Linear + activation (base_modules/linear.py)"""

    def __init__(self, input_dim, out_dim, activation=None, *args, **kwargs):
        super().__init__(input_dim, out_dim, *args, **kwargs)
        self.activation = nn.ReLU() if activation is None else activation

    def forward(self, x):
        return self.activation(super().forward(x))


@MODULE_REGISTRY.register()
class GeoNet(nn.Module):
    def __init__(self, W=256, D=8, skips=[4], encoder=None, W_feat=256, use_bias=True, skip_reduce_output=False,
                 norm_skip=False, act_cfg=None, geometric_init=True, radius_init=1.0, use_siren=False, weight_norm=False,
                 out_act_cfg=None, *args, **kwargs):
        super(GeoNet, self).__init__()
        if use_siren:
            raise NotImplementedError('Siren layers are not mirrored')
        self.W, self.D, self.skips, self.W_feat, self.norm_skip = W, D, list(skips), W_feat, norm_skip
        self.out_act = get_activation(out_act_cfg) if out_act_cfg is not None else None
        self.embed_fn, input_ch, embed_freq = build_encoder(encoder)
        self.embed_dim = self.embed_fn.get_output_dim()
        layers = []
        for i in range(D + 1):
            if i == 0:
                in_dim = self.embed_dim
            elif not skip_reduce_output and (i - 1) in self.skips:
                in_dim = self.embed_dim + W
            else:
                in_dim = W
            if i == D:
                out_dim = 1 + W_feat if W_feat > 0 else 1
            elif skip_reduce_output and i in self.skips:
                out_dim = W - self.embed_dim
            else:
                out_dim = W
            layer = DenseLayer(in_dim, out_dim, activation=get_activation(act_cfg), bias=use_bias) if i != D \
                else Linear(in_dim, out_dim, bias=use_bias)
            if geometric_init:                                    # linear_network_module.py:134-157
                if i == D:
                    nn.init.normal_(layer.weight, mean=np.sqrt(np.pi) / np.sqrt(in_dim), std=0.0001)
                    if use_bias:
                        nn.init.constant_(layer.bias[:1], -radius_init)
                elif embed_freq > 0:
                    if use_bias:
                        nn.init.constant_(layer.bias, 0.0)
                    if i == 0:
                        nn.init.constant_(layer.weight[:, input_ch:], 0.0)
                        nn.init.normal_(layer.weight[:, :input_ch], 0.0, np.sqrt(2) / np.sqrt(out_dim))
                    elif (i - 1) in self.skips:
                        nn.init.normal_(layer.weight, 0.0, np.sqrt(2) / np.sqrt(out_dim))
                        nn.init.constant_(layer.weight[:, -(self.embed_dim - input_ch):], 0.0)
                    else:
                        nn.init.normal_(layer.weight, 0.0, np.sqrt(2) / np.sqrt(out_dim))
                else:
                    if use_bias:
                        nn.init.constant_(layer.bias, 0.0)
                    nn.init.normal_(layer.weight, 0.0, np.sqrt(2) / np.sqrt(out_dim))
            if weight_norm:
                layer = nn.utils.weight_norm(layer)
            layers.append(layer)
        self.layers = nn.ModuleList(layers)

    def handle_output(self, out):
        if self.W_feat <= 0:
            out_geo, out_feat = out, None
        else:
            out_geo, out_feat = out[:, 0].unsqueeze(-1), out[:, 1:]
        if self.out_act is not None:
            out_geo = self.out_act(out_geo)
        return out_geo, out_feat

    def forward(self, x):
        """(B, input_ch) -> geo (B, 1), feat (B, W_feat) | None"""
        x_embed = self.embed_fn(x)
        out = x_embed
        for i in range(self.D + 1):
            out = self.layers[i](out)
            if i in self.skips:
                out = torch.cat([out, x_embed], dim=-1)
                if self.norm_skip:
                    out = out / math.sqrt(2)
        return self.handle_output(out)

    def forward_geo_value(self, x):
        return self.forward(x)[0][:, 0]

    def forward_with_grad(self, x):
        """base_network.py:30-44 -> geo, feat, d geo / d x (kept in the graph for the eikonal term)"""
        with torch.enable_grad():
            x = x.requires_grad_(True)
            geo, feat = self.forward(x)
            grad = torch.autograd.grad(geo, x, torch.ones_like(geo), create_graph=True, retain_graph=True, only_inputs=True)[0]
        return geo, feat, grad


@MODULE_REGISTRY.register()
class RadianceNet(nn.Module):
    def __init__(self, mode='vf', W=256, D=8, encoder=None, W_feat_in=256, use_bias=True, act_cfg=None, use_siren=False,
                 weight_norm=False, out_act_cfg=None, *args, **kwargs):
        super(RadianceNet, self).__init__()
        if use_siren:
            raise NotImplementedError('Siren layers are not mirrored')
        assert len(mode) > 0 and all([m in 'pvnf' for m in mode]), 'Invalid mode only pvnf allowed...'
        self.mode, self.W, self.D, self.W_feat_in = mode, W, D, W_feat_in
        enc = cfg_to_dict(encoder)
        self.init_input_dim = 0
        self.embed_fn_pts = self.embed_fn_view = None
        if 'p' in mode:
            self.embed_fn_pts, _, _ = build_encoder(enc.get('pts'))
            self.init_input_dim += self.embed_fn_pts.get_output_dim()
        if 'v' in mode:
            self.embed_fn_view, _, _ = build_encoder(enc.get('view'))
            self.init_input_dim += self.embed_fn_view.get_output_dim()
        if 'n' in mode:
            self.init_input_dim += 3
        if 'f' in mode and W_feat_in > 0:
            self.init_input_dim += W_feat_in
        layers = []
        for i in range(D + 1):
            in_dim = self.init_input_dim if i == 0 else W
            if i != D:
                layer = DenseLayer(in_dim, W, activation=get_activation(act_cfg), bias=use_bias)
            else:
                layer = DenseLayer(in_dim, 3, activation=get_activation(out_act_cfg, 'sigmoid'), bias=use_bias)
            if weight_norm:
                layer = nn.utils.weight_norm(layer)
            layers.append(layer)
        self.layers = nn.ModuleList(layers)

    def fuse_radiance_inputs(self, x, view_dirs, normals, geo_feat):
        """encoder_mlp_network.py:93-118: collected in p, v, n, f order, then permuted by the argsort of the mode letters' ranks
        exactly as the reference does (for two-letter modes: the order of the mode string)"""
        inputs = {}
        if 'p' in self.mode:
            inputs['p'] = self.embed_fn_pts(x)
        if 'v' in self.mode:
            inputs['v'] = self.embed_fn_view(view_dirs / torch.norm(view_dirs, dim=-1, keepdim=True))
        if 'n' in self.mode:
            inputs['n'] = normals
        if 'f' in self.mode:
            inputs['f'] = geo_feat
        parts = [inputs[k] for k in 'pvnf' if k in inputs]
        rank = {'p': 0, 'v': 1, 'n': 2, 'f': 3}
        order = np.argsort([rank[m] for m in self.mode]).tolist()
        out = torch.cat([parts[i] for i in order], dim=-1)
        assert out.shape[-1] == self.init_input_dim, 'Shape not match'
        return out

    def forward(self, x, view_dirs, normals, geo_feat):
        out = self.fuse_radiance_inputs(x, view_dirs, normals, geo_feat)
        for i in range(self.D + 1):
            out = self.layers[i](out)
        return out
