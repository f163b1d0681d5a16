# -*- coding: utf-8 -*-
"""Drop-in for arcnerf/models/base_modules/geo_rad_model/tcnn_fusedmlp_module.py: FusedMLPGeoNet and
FusedMLPRadianceNet on the libarcb200 tensor-core MLP instead of tinycudann.Network.

Same constructor arguments and forward contracts (tcnn_fusedmlp_module.py:16-192):
  FusedMLPGeoNet.forward(x)                       -> (geo (B,1) [after out_act], feat (B,W_feat) or None)
  FusedMLPRadianceNet.forward(x, view, normal, f) -> (B,3)
The network parameters live in `layers.params`, flat fp32 in tiny-cuda-nn's layout (same state_dict key and layout
as the reference's `tcnn.Network`).
"""
import math

import torch
import torch.nn as nn
import torch.nn.functional as tnf

from .. import functional as F
from .registry import MODULE_REGISTRY, build_encoder, cfg_to_dict
from .trunc_exp import TruncExp


def get_tcnn_activation_from_cfgs(cfg, default='None'):
    d = cfg_to_dict(cfg)
    if not d:
        return default
    t = d['type'].lower()
    return {'relu': 'ReLU', 'softplus': 'Softplus', 'sigmoid': 'Sigmoid', 'exponential': 'Exponential', 'none': 'None'}[t]


def get_activation(cfg):
    """Output activations ArcNerf applies in torch on top of the network (base_modules/activation.py): TruncExp, ReLU,
    Softplus, Sigmoid."""
    d = cfg_to_dict(cfg)
    if not d:
        return None
    t = d['type'].lower()
    if t == 'truncexp':
        return TruncExp(d.get('clip', 15.0))
    if t == 'relu':
        return nn.ReLU()
    if t == 'softplus':
        return nn.Softplus(beta=d.get('beta', 1))
    if t == 'sigmoid':
        return nn.Sigmoid()
    raise NotImplementedError('activation %s' % d['type'])


class _MlpFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mod, x, params):
        ctx.mod = mod
        xc = x.contiguous()
        if xc.dtype not in (torch.float16, torch.float32):
            xc = xc.float()
        ctx.save_for_backward(xc, params)
        ctx.x_dtype, ctx.need_dx = x.dtype, x.requires_grad
        return F.mlp_forward(xc, mod.blob(), mod.spec, out_dtype=torch.float16)

    @staticmethod
    def backward(ctx, dy):
        xc, params = ctx.saved_tensors
        mod = ctx.mod
        dx, dw = F.mlp_backward(xc, mod.blob(), mod.spec, dy.contiguous().float(), want_dx=ctx.need_dx,
                                loss_scale=mod.loss_scale)
        return None, (dx.to(ctx.x_dtype) if ctx.need_dx else None), dw


class FusedMLP(nn.Module):
    """tinycudann.Network(otype=FullyFusedMLP) stand-in: flat fp32 `params`, fp16 output (B, n_output_dims)."""

    def __init__(self, n_input_dims, n_output_dims, network_config, seed=1337):
        super().__init__()
        assert network_config.get('otype', 'FullyFusedMLP') == 'FullyFusedMLP'
        self.n_input_dims, self.n_output_dims = n_input_dims, n_output_dims
        self.spec = F.MlpSpec(n_input_dims, network_config['n_neurons'], network_config['n_hidden_layers'], n_output_dims,
                              network_config.get('activation', 'ReLU'), network_config.get('output_activation', 'None'))
        g = torch.Generator().manual_seed(seed)
        w, in_pad, out_pad = self.spec.width, self.spec.in_pad, self.spec.out_pad
        shapes = [(w, in_pad)] + [(w, w)] * (self.spec.n_hidden - 1) + [(out_pad, w)]
        mats = [(torch.rand(o * i, generator=g) * 2 - 1) * math.sqrt(6.0 / (i + o)) for o, i in shapes]   # xavier uniform
        self.params = nn.Parameter(torch.cat(mats))
        self.loss_scale = 128.0                      # tinycudann/modules.py:141
        self._blob, self._blob_key = None, None

    def blob(self):
        key = (self.params.data_ptr(), self.params._version)
        if self._blob is None or self._blob_key != key or self._blob.device != self.params.device:
            self._blob = F.mlp_pack(self.params.detach().contiguous(), self.spec,
                                    self._blob if self._blob is not None and self._blob.device == self.params.device else None)
            self._blob_key = key
        return self._blob

    def forward(self, x):
        return _MlpFn.apply(self, x, self.params)


@MODULE_REGISTRY.register()
class FusedMLPGeoNet(nn.Module):
    def __init__(self, W=128, D=8, encoder=None, W_feat=128, act_cfg=None, out_act_cfg=None, dtype=torch.float32,
                 input_ch=3, *args, **kwargs):
        super(FusedMLPGeoNet, self).__init__()
        self.W, self.D, self.W_feat, self.dtype = W, D, W_feat, dtype
        self.out_act = get_activation(out_act_cfg)
        self.embed_fn, embed_dim = build_encoder(encoder)
        self.embed_dim = embed_dim if embed_dim is not None else input_ch
        if W_feat > 0:
            assert W_feat in [8, 16, 32, 64, 128], 'Restrict num of layers for fused mlp.'
        self.layers = FusedMLP(self.embed_dim, W_feat if W_feat > 0 else 1, {
            'otype': 'FullyFusedMLP', 'activation': get_tcnn_activation_from_cfgs(act_cfg, 'ReLU'),
            'output_activation': 'None', 'n_neurons': W, 'n_hidden_layers': D})

    def forward(self, x):
        x_embed = self.embed_fn(x) if self.embed_fn is not None else x
        out = self.layers(x_embed).type(self.dtype)
        return self.handle_output_combine(out)

    def handle_output_combine(self, out):
        """first output channel is the geometry value, all W_feat channels are the feature (tcnn_fusedmlp_module.py:99-111)"""
        if self.W_feat <= 0:
            out_geo, out_feat = out, None
        else:
            out_geo, out_feat = out[:, 0].unsqueeze(-1), out
        if self.out_act is not None:
            out_geo = self.out_act(out_geo)
        return out_geo, out_feat

    def forward_geo_value(self, x):
        return self.forward(x)[0][:, 0]

    def forward_with_grad(self, x):
        """geo value, feature and d(geo)/dx by autograd (base_network.py:30-44)"""
        with torch.enable_grad():
            x = x.requires_grad_(True)
            geo, h = self.forward(x)
            grad = torch.autograd.grad(outputs=geo, inputs=x, grad_outputs=torch.ones_like(geo), create_graph=True,
                                       retain_graph=True, only_inputs=True)[0]
        return geo, h, grad


@MODULE_REGISTRY.register()
class FusedMLPRadianceNet(nn.Module):
    def __init__(self, mode='vf', W=128, D=8, encoder=None, W_feat_in=128, act_cfg=None, out_act_cfg=None,
                 dtype=torch.float32, input_ch_pts=3, input_ch_view=3, *args, **kwargs):
        super(FusedMLPRadianceNet, self).__init__()
        assert len(mode) > 0 and all(m in 'pvnf' for m in mode), 'Invalid mode only pvnf allowed...'
        self.mode, self.W, self.D, self.W_feat_in, self.dtype = mode, W, D, W_feat_in, dtype
        enc = cfg_to_dict(encoder)
        self.embed_fn_pts, pts_dim = build_encoder(enc.get('pts')) if 'p' in mode else (None, None)
        self.embed_fn_view, view_dim = build_encoder(enc.get('view')) if 'v' in mode else (None, None)
        self.init_input_dim = 0
        if 'p' in mode:
            self.init_input_dim += pts_dim if pts_dim is not None else input_ch_pts
        if 'v' in mode:
            self.init_input_dim += view_dim if view_dim is not None else input_ch_view
        if 'n' in mode:
            self.init_input_dim += 3
        if 'f' in mode and W_feat_in > 0:
            self.init_input_dim += W_feat_in
        self.layers = FusedMLP(self.init_input_dim, 3, {
            'otype': 'FullyFusedMLP', 'activation': get_tcnn_activation_from_cfgs(act_cfg, 'ReLU'),
            'output_activation': get_tcnn_activation_from_cfgs(out_act_cfg, 'Sigmoid'), 'n_neurons': W,
            'n_hidden_layers': D})

    def fuse_radiance_inputs(self, x, view_dirs, normals, geo_feat):
        """collect in p,v,n,f order, then permute by argsort of the mode letters' ranks exactly as the reference does
        (encoder_mlp_network.py:93-118; for two-letter modes this is the order of the mode string, e.g. 'fv' ->
        [feat, view])"""
        parts = {}
        if 'p' in self.mode:
            parts['p'] = self.embed_fn_pts(x) if self.embed_fn_pts is not None else x
        if 'v' in self.mode:
            v = tnf.normalize(view_dirs, dim=-1)
            parts['v'] = self.embed_fn_view(v) if self.embed_fn_view is not None else v
        if 'n' in self.mode:
            parts['n'] = normals
        if 'f' in self.mode:
            parts['f'] = geo_feat
        inputs = [parts[k] for k in 'pvnf' if k in parts]
        rank = {'p': 0, 'v': 1, 'n': 2, 'f': 3}
        ranks = [rank[m] for m in self.mode]
        order = sorted(range(len(ranks)), key=lambda i: ranks[i])          # np.argsort
        out = torch.cat([inputs[i].float() for i in order], dim=-1)
        assert out.shape[-1] == self.init_input_dim, 'Shape not match'
        return out

    def forward(self, x, view_dirs, normals, geo_feat):
        out = self.fuse_radiance_inputs(x, view_dirs, normals, geo_feat)
        return self.layers(out).type(self.dtype)
