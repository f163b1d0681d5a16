# -*- coding: utf-8 -*-
"""Drop-in for arcnerf/models/base_modules/encoding/freq_encoder.py:FreqEmbedder on libarcb200's
frequency-encoding kernels: same constructor arguments, `get_output_dim()`, `forward((B, input_dim)) -> (B, out_dim)`,
differentiable w.r.t. the input (surface normals of SDF models go through it).  Periodic functions are (sin, cos)."""
import ctypes as C

import torch
import torch.nn as nn

from .. import _lib
from .registry import ENCODER_REGISTRY


class _FreqFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, freqs_c, n_freqs, include_input, out_dim):
        xc = x.contiguous().float()
        out = torch.empty((xc.shape[0], out_dim), dtype=torch.float32, device=xc.device)
        _lib.call('arcb200_freq_embed_fwd', xc.data_ptr(), xc.shape[1], freqs_c, n_freqs, int(include_input), out.data_ptr(),
                  out_dim, xc.shape[0], _lib.current_stream())
        ctx.save_for_backward(xc)
        ctx.args = (freqs_c, n_freqs, include_input, out_dim)
        return out

    @staticmethod
    def backward(ctx, g):
        (xc, ) = ctx.saved_tensors
        return _FreqBwdFn.apply(g, xc, ctx.args), None, None, None, None


class _FreqBwdFn(torch.autograd.Function):
    """The input gradient as a function of (g, x) that can be differentiated again: the normals of an SDF network behind a
    frequency embedding enter the eikonal loss, whose backward needs d(g_x)/d(g) and d(g_x)/d(x)."""

    @staticmethod
    def forward(ctx, g, xc, args):
        freqs_c, n_freqs, include_input, out_dim = args
        gc = g.contiguous().float()
        gx = torch.empty_like(xc)
        _lib.call('arcb200_freq_embed_bwd', xc.data_ptr(), xc.shape[1], freqs_c, n_freqs, int(include_input), gc.data_ptr(),
                  out_dim, gx.data_ptr(), xc.shape[0], _lib.current_stream())
        ctx.save_for_backward(gc, xc)
        ctx.args = args
        return gx

    @staticmethod
    def backward(ctx, v):
        """g_x = g_0 + sum_k f_k (cos(x f_k) g_sin,k - sin(x f_k) g_cos,k): linear in g (Jacobian of the embedding applied to v)
        and a second derivative in x.  Second-order passes are rare and small: elementwise torch expressions."""
        gc, xc = ctx.saved_tensors
        freqs_c, n_freqs, include_input, out_dim = ctx.args
        d = xc.shape[1]
        parts, dx = [], torch.zeros_like(xc)
        col = 0
        if include_input:
            parts.append(v)
            col = d
        for k in range(n_freqs):
            f = float(freqs_c[k])
            s, c = torch.sin(xc * f), torch.cos(xc * f)
            parts += [v * f * c, -v * f * s]
            dx = dx + v * (f * f) * (-s * gc[:, col:col + d] - c * gc[:, col + d:col + 2 * d])
            col += 2 * d
        return torch.cat(parts, 1), dx, None


@ENCODER_REGISTRY.register()
class FreqEmbedder(nn.Module):
    """freq_encoder.py:10-88"""

    def __init__(self, input_dim, n_freqs, log_sampling=True, include_input=True, periodic_fns=None, *args, **kwargs):
        super(FreqEmbedder, self).__init__()
        if periodic_fns is not None and tuple(periodic_fns) != (torch.sin, torch.cos):
            raise NotImplementedError('the kernel evaluates (sin, cos), the reference default')
        self.input_dim = input_dim
        self.include_input = include_input
        self.out_dim = (input_dim if include_input else 0) + input_dim * n_freqs * 2
        if n_freqs == 0 and include_input:
            self.freq_bands = []
        elif log_sampling:
            self.freq_bands = 2.**torch.linspace(0., n_freqs - 1, n_freqs)
        else:
            self.freq_bands = torch.linspace(2.**0., 2.**(n_freqs - 1), n_freqs)
        bands = [float(f) for f in self.freq_bands]
        self._n = len(bands)
        self._freqs_c = (C.c_float * max(1, self._n))(*bands)

    def get_output_dim(self):
        return self.out_dim

    def forward(self, x):
        assert (x.shape[-1] == self.input_dim), 'Input shape should be (B, {})'.format(self.input_dim)
        if not x.is_cuda:
            raise RuntimeError('x must be a CUDA tensor')
        return _FreqFn.apply(x, self._freqs_c, self._n, self.include_input, self.out_dim)
