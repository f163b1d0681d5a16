# -*- coding: utf-8 -*-
"""The vanilla-NeRF geometry + radiance networks as one tensor-core chain (csrc/dense_chain.cu).

`FusedNeRFNets` stands where the reference builds `GeoNet(W=256, D=8, skips=[4], FreqEmbedder(10))` and
`RadianceNet(mode='vf', W=128, D=1, FreqEmbedder(4))` (arcnerf/models/base_modules/geo_rad_model/linear_network_module.py:
17-197,234-335; configs/models/nerf.yaml, nerfpp.yaml) and calls them back to back on the same points
(Base3dModel._forward_pts_dir, base_3d_model.py:233-254).  Same parameters (the reference's `layers.{i}.weight / bias`
tensors, exposed as views of one flat vector so that its state dicts load), same outputs, gradients w.r.t. every parameter
through the library's backward kernels.  No torch.nn.Linear, no cuBLAS."""
import ctypes as C
import math

import numpy as np
import torch
import torch.nn as nn

from .. import _lib


class _ChainFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, nets, x, dirs, flat, need_grad):
        n = x.shape[0]
        sigma = torch.empty((n,), dtype=torch.float32, device=x.device)
        rgb = torch.empty((n, 3), dtype=torch.float32, device=x.device)
        nets._pack_if_changed()
        n_cap = (n + 127) // 128 * 128
        save = torch.empty(nets.fwd_cols * n_cap, dtype=torch.float16, device=x.device) if need_grad else None
        _lib.call('arcb200_nerf_mlp_fwd', *nets.arch, x.data_ptr(), dirs.data_ptr(), nets.blob_fwd.data_ptr(), nets.vec.data_ptr(),
                  _lib.ptr(save), n_cap, sigma.data_ptr(), rgb.data_ptr(), n, _lib.current_stream())
        ctx.nets, ctx.n, ctx.n_cap = nets, n, n_cap
        if nets.keep_save:
            nets.last_save = (save, n_cap)
        ctx.save_for_backward(save, rgb)
        return sigma, rgb

    @staticmethod
    def backward(ctx, g_sigma, g_rgb):
        nets, n, n_cap = ctx.nets, ctx.n, ctx.n_cap
        save, rgb = ctx.saved_tensors
        dev = rgb.device
        g_sigma = torch.zeros((n,), device=dev) if g_sigma is None else g_sigma.contiguous().float()
        g_rgb = torch.zeros((n, 3), device=dev) if g_rgb is None else g_rgb.contiguous().float()
        bwd_save = torch.empty(nets.bwd_cols * n_cap, dtype=torch.float16, device=dev)
        dsig16 = torch.empty((n_cap, 16), dtype=torch.float16, device=dev)
        d_params = torch.zeros(nets.n_params, dtype=torch.float32, device=dev)
        _lib.call('arcb200_nerf_mlp_bwd', *nets.arch, nets.blob_bwd.data_ptr(), nets.vec.data_ptr(), save.data_ptr(), rgb.data_ptr(),
                  g_sigma.data_ptr(), g_rgb.data_ptr(), float(nets.loss_scale), bwd_save.data_ptr(), dsig16.data_ptr(), n_cap, d_params.data_ptr(), n,
                  _lib.current_stream())
        return None, None, None, d_params, None


class FusedNeRFNets(nn.Module):
    def __init__(self, d_x=3, n_freq_x=10, n_freq_v=4, W=256, D=8, skips=(4,), W_rad=128, device='cuda:0'):
        super().__init__()
        skips = list(skips)
        assert len(skips) <= 1, 'one skip connection or none'
        self.arch = (int(d_x), int(n_freq_x), int(n_freq_v), int(W), int(D), int(skips[0]) if skips else -1, int(W_rad))
        sizes = (C.c_int64 * 6)()
        _lib.call('arcb200_nerf_mlp_sizes', *self.arch, sizes)
        self.n_params, fwd_bytes, bwd_bytes, n_vec, self.fwd_cols, self.bwd_cols = [int(v) for v in sizes]
        dev = torch.device(device)
        self.embed_x, self.embed_v = d_x * (2 * n_freq_x + 1), 3 * (2 * n_freq_v + 1)
        # (out, in) of every linear layer in parameter order
        self.shapes = []
        for i in range(D + 1):
            fan_in = self.embed_x if i == 0 else (W + self.embed_x if (skips and i - 1 == skips[0]) else W)
            self.shapes.append((1 + W if i == D else W, fan_in))
        self.shapes += [(W_rad, self.embed_v + W), (3, W_rad)]
        flat = torch.empty(self.n_params, dtype=torch.float32)
        off = 0
        for (o, i) in self.shapes:          # torch.nn.Linear's default initialisation (what the reference's DenseLayer keeps)
            bound = 1.0 / math.sqrt(i)
            flat[off:off + o * i + o].uniform_(-bound, bound)
            off += o * i + o
        assert off == self.n_params
        self.flat = nn.Parameter(flat.to(dev))
        self.blob_fwd = torch.empty(fwd_bytes, dtype=torch.uint8, device=dev)
        self.blob_bwd = torch.empty(bwd_bytes, dtype=torch.uint8, device=dev)
        self.vec = torch.zeros(n_vec, dtype=torch.float32, device=dev)
        self._packed_version = None
        self.keep_save, self.last_save = False, None     # tests: keep the forward's saved activations for inspection
        self.loss_scale = 128.0        # gradients cross the fp16 tiles multiplied by this (tinycudann/modules.py:141 uses the same)

    # ---- the reference's parameter tensors as views of the flat vector
    def layer_views(self):
        """[(weight (out,in), bias (out,))] for geometry layers 0..D, then the two radiance layers"""
        out, off = [], 0
        for (o, i) in self.shapes:
            out.append((self.flat.data[off:off + o * i].view(o, i), self.flat.data[off + o * i:off + o * i + o]))
            off += o * i + o
        return out

    def load_reference_state(self, geo_state, rad_state):
        """state dicts of the reference's GeoNet / RadianceNet (`layers.{i}.weight`, `layers.{i}.bias`)"""
        views = self.layer_views()
        n_geo = self.arch[4] + 1
        with torch.no_grad():
            for i in range(n_geo):
                views[i][0].copy_(torch.as_tensor(geo_state['layers.%d.weight' % i]))
                views[i][1].copy_(torch.as_tensor(geo_state['layers.%d.bias' % i]))
            for i in range(2):
                views[n_geo + i][0].copy_(torch.as_tensor(rad_state['layers.%d.weight' % i]))
                views[n_geo + i][1].copy_(torch.as_tensor(rad_state['layers.%d.bias' % i]))
        self._packed_version = None

    def reference_state(self):
        views = self.layer_views()
        n_geo = self.arch[4] + 1
        geo = {}
        for i in range(n_geo):
            geo['layers.%d.weight' % i], geo['layers.%d.bias' % i] = views[i][0].clone(), views[i][1].clone()
        rad = {}
        for i in range(2):
            rad['layers.%d.weight' % i], rad['layers.%d.bias' % i] = views[n_geo + i][0].clone(), views[n_geo + i][1].clone()
        return geo, rad

    @staticmethod
    def untile(region, n_cap, width):
        """a saved tensor (flat fp16 region of n_cap * width elements in the operand-tile layout of csrc/dense_chain.cu: 128-row
        tiles, inside a tile 8-column groups of 128 rows) -> (n_cap, width) row-major"""
        return region.view(n_cap // 128, width // 8, 128, 8).permute(0, 2, 1, 3).reshape(n_cap, width)

    def saved_columns(self):
        """name -> (first column, width) of the forward save buffer (region c starts at c * n_cap and holds n_cap * width
        elements in the tiled layout, see untile)"""
        d_x, nfx, nfv, W, D, skip, W_rad = self.arch
        xe_pad = (self.embed_x + 31) // 32 * 32
        cols, c = {'xe': (0, xe_pad), 've': (xe_pad, 32)}, xe_pad + 32
        for i in range(D):
            cols['h%d' % i] = (c, W); c += W
        cols['feat'] = (c, W); c += W
        cols['hr'] = (c, W_rad); c += W_rad
        assert c == self.fwd_cols
        return cols

    def _pack_if_changed(self):
        v = self.flat._version
        if self._packed_version != v:
            _lib.call('arcb200_nerf_mlp_pack', *self.arch, self.flat.data_ptr(), self.blob_fwd.data_ptr(), self.blob_bwd.data_ptr(),
                      self.vec.data_ptr(), _lib.current_stream())
            self._packed_version = v

    def forward(self, x, view_dirs):
        """x (B, d_x), view_dirs (B, 3) -> sigma (B, 1), rgb (B, 3)   (GeoNet.forward followed by RadianceNet.forward)"""
        if not x.is_cuda or not view_dirs.is_cuda:
            raise RuntimeError('FusedNeRFNets needs CUDA tensors')
        assert x.dim() == 2 and x.shape[1] == self.arch[0] and view_dirs.shape == (x.shape[0], 3), 'x (B,d_x), view_dirs (B,3)'
        # activations are kept for a backward only when one can follow (grad mode is not visible inside Function.forward)
        need_grad = torch.is_grad_enabled() and self.flat.requires_grad
        sigma, rgb = _ChainFn.apply(self, x.contiguous().float(), view_dirs.contiguous().float(), self.flat, need_grad)
        return sigma.unsqueeze(-1), rgb
