# -*- coding: utf-8 -*-
"""Drop-in for arcnerf/models/base_modules/encoding/hashgrid_encoder.py with `backend: b200` (or `tcnn`, which here
means the same kernels): the multiresolution hash-grid embedder, tiny-cuda-nn semantics, on libarcb200.

Same constructor arguments, `get_output_dim()`, `forward((B,3)) -> (B, L*F [+3])`, and the same state_dict key for the
table: `hashgrid_encode_tcnn.params`, a flat fp32 tensor in tiny-cuda-nn's layout, so checkpoints written by the
reference load here and vice versa (SURVEY.md §5 checkpoint row).  First and second derivatives w.r.t. the input are
supported (NeuS normals + eikonal loss: base_network.py:30-44, tinycudann/modules.py:73-127).
"""
import torch
import torch.nn as nn

from .. import functional as F
from .registry import ENCODER_REGISTRY


def cast_torch_dtype(dtype):
    if isinstance(dtype, torch.dtype):
        return dtype
    return {'torch.float16': torch.float16, 'torch.float32': torch.float32, 'torch.float64': torch.float64}[dtype]


class _GridFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mod, x, params):
        need_dx = x.requires_grad
        th = mod.table_half()
        if need_dx:
            y, dydx = F.grid_forward(x, th, mod.geom, want_dydx=True)
        else:
            y, dydx = F.grid_forward(x, th, mod.geom), None
        ctx.mod, ctx.need_dx = mod, need_dx
        ctx.save_for_backward(x, params, dydx if dydx is not None else torch.empty(0, device=x.device))
        return y

    @staticmethod
    def backward(ctx, dy):
        x, params, dydx = ctx.saved_tensors
        dx, dp = _GridBwdFn.apply(ctx.mod, dy.contiguous(), x, params, dydx, ctx.need_dx)
        return None, (dx if ctx.need_dx else None), dp


class _GridBwdFn(torch.autograd.Function):
    """The backward pass as a differentiable function, so that dL/dx can itself be differentiated
    (tinycudann/modules.py:99-127, _module_function_backward)."""

    @staticmethod
    def forward(ctx, mod, dy, x, params, dydx, need_dx):
        ctx.mod = mod
        ctx.save_for_backward(dy, x, params)
        dp = F.grid_backward(x, dy, mod.geom)
        dx = F.grid_backward_input(dy, dydx, mod.geom) if need_dx else torch.zeros_like(x)
        return dx, dp

    @staticmethod
    def backward(ctx, d_dx, d_dp):
        # d_dp (second order through the parameter gradient) is unsupported, as in the reference
        dy, x, params = ctx.saved_tensors
        mod = ctx.mod
        if d_dx is None:
            return None, None, None, None, None, None
        gt = torch.zeros(mod.geom.n_params, dtype=torch.float32, device=x.device)
        gt, d_dy, d_x = F.grid_double_backward(x, mod.table_half(), dy, d_dx.contiguous().float(), mod.geom, grad_table=gt)
        return None, d_dy.to(dy.dtype), d_x, gt, None, None


class _GridModule(nn.Module):
    """Holds the flat fp32 `params` (tcnn layout) and its cached fp16 working copy."""

    def __init__(self, geom, seed=1337):
        super().__init__()
        self.geom = geom
        g = torch.Generator().manual_seed(seed)
        init = (torch.rand(geom.n_params, generator=g) * 2 - 1) * 1e-4      # tcnn grid.h:1353-1356
        self.params = nn.Parameter(init)
        self._half, self._half_key = None, None

    def table_half(self):
        key = (self.params.data_ptr(), self.params._version)
        if self._half is None or self._half_key != key or self._half.device != self.params.device:
            self._half = F.cast_f16(self.params.detach().contiguous(), self._half if self._half is not None and
                                    self._half.device == self.params.device else None)
            self._half_key = key
        return self._half

    def forward(self, x):
        return _GridFn.apply(self, x.contiguous().float(), self.params)


@ENCODER_REGISTRY.register()
class HashGridEmbedder(nn.Module):
    def __init__(self, input_dim=3, n_levels=16, n_feat_per_entry=2, hashmap_size=19, base_res=16, max_res=2048,
                 origin=(0, 0, 0), side=None, xyz_len=None, dtype='torch.float16', include_input=True, backend=None,
                 *args, **kwargs):
        super(HashGridEmbedder, self).__init__()
        assert input_dim == 3, 'HashGridEmbedder should has input_dim==3...'
        assert side is not None or xyz_len is not None, 'You must set the size of volume...'
        if backend is None:
            backend = 'b200'
        assert backend in ['b200', 'tcnn'], 'Invalid backend: this build provides the CUDA (b200/tcnn-compatible) path only'
        self.backend = backend
        self.input_dim, self.include_input = input_dim, include_input
        self.dtype = cast_torch_dtype(dtype)
        self.base_res, self.max_res = base_res, max_res
        self.per_level_scale = torch.exp((torch.log(torch.tensor(max_res / base_res))) / (float(n_levels) - 1))
        self.hashmap_size, self.n_feat_per_entry, self.n_levels = 2 ** hashmap_size, n_feat_per_entry, n_levels
        geom = F.GridGeometry(n_levels, n_feat_per_entry, hashmap_size, base_res, float(self.per_level_scale))
        self.hashgrid_encode_tcnn = _GridModule(geom)
        # volume range (Volume(...).get_range(), geometry/volume.py): centre +- half side
        origin = torch.tensor(origin, dtype=torch.float32)
        half = torch.tensor([side / 2.0] * 3 if side is not None else [v / 2.0 for v in xyz_len], dtype=torch.float32)
        self.register_buffer('min_xyz', origin - half)
        self.register_buffer('max_xyz', origin + half)
        self.out_dim = n_levels * n_feat_per_entry + include_input * input_dim

    def get_output_dim(self):
        return self.out_dim

    def forward(self, xyz):
        assert len(xyz.shape) == 2 and xyz.shape[-1] == 3, 'Must be (B, 3) tensor'
        out = []
        if self.include_input:
            out.append(xyz)
        norm_xyz = (xyz - self.min_xyz) / (self.max_xyz - self.min_xyz)
        out.append(self.hashgrid_encode_tcnn(norm_xyz).type(self.dtype))
        return torch.cat(out, dim=-1) if len(out) > 1 else out[0]
