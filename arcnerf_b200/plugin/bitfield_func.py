# -*- coding: utf-8 -*-
"""Drop-in for arcnerf/ops/bitfield_func.py (Morton-bitfield sampler and density-grid maintenance),
backed by libarcb200 instead of the `_bitfield_func` pybind module (bitfield_func.py:17-159)."""
import torch

from .. import _lib
from ._common import check, f32c
from .volume_func import _sampling_args

CUDA_BACKEND_AVAILABLE = True

rng = None   # per-module generator, as in the reference (include/common.h:21-23)


def _rng():
    global rng
    if rng is None:
        rng = _lib.Rng()
    return rng


def reset_rng(seed=_lib.Rng.SEED):
    global rng
    rng = _lib.Rng(seed)


@torch.no_grad()
def sparse_volume_sampling_bit(rays_o, rays_d, near, far, n_pts, dt, aabb_range, n_grid, bitfield, near_distance=0.0):
    """As volume_func.sparse_volume_sampling with a (n_grid**3 // 8,) uint8 Morton bitfield
    (bitfield_func.py:37-61)."""
    rays_o, rays_d, near, far, aabb, n_rays = _sampling_args(rays_o, rays_d, near, far, aabb_range)
    check(bitfield, 'bitfield', torch.uint8)
    if bitfield.numel() != n_grid ** 3 // 8:
        raise RuntimeError('bitfield must hold n_grid**3 / 8 bytes')
    zvals = torch.empty((n_rays, n_pts), dtype=torch.float32, device=rays_o.device)
    mask = torch.empty((n_rays, n_pts), dtype=torch.bool, device=rays_o.device)
    r = _rng()
    _lib.call('arcb200_sparse_volume_sampling_bit', rays_o.data_ptr(), rays_d.data_ptr(), near.data_ptr(),
              far.data_ptr(), int(n_pts), float(dt), aabb.data_ptr(), int(n_grid), bitfield.data_ptr(),
              float(near_distance), zvals.data_ptr(), mask.data_ptr(), n_rays, r.state, r.inc, _lib.current_stream())
    r.advance()
    return zvals, mask


@torch.no_grad()
def generate_grid_samples(density_grid, n_elements, density_grid_ema_step, n_grid, thresh):
    """-> positions (n_elements,3) in [0,1), indices (n_elements,) int32   (bitfield_func.py:67-88)"""
    density_grid = check(density_grid, 'density_grid', torch.float32)
    dev = density_grid.device
    positions = torch.empty((n_elements, 3), dtype=torch.float32, device=dev)
    indices = torch.empty((n_elements, ), dtype=torch.int32, device=dev)
    r = _rng()
    _lib.call('arcb200_generate_grid_samples', density_grid.data_ptr(), int(density_grid_ema_step), int(n_elements),
              int(n_grid), float(thresh), positions.data_ptr(), indices.data_ptr(), r.state, r.inc,
              _lib.current_stream())
    r.advance()
    return positions, indices


@torch.no_grad()
def splat_grid_samples(density, density_grid_indices, n_samples, density_grid_tmp):
    """density_grid_tmp[idx] = max(density_grid_tmp[idx], density)   (bitfield_func.py:94-106)"""
    density = check(f32c(density), 'density', torch.float32)
    check(density_grid_indices, 'density_grid_indices', torch.int32)
    check(density_grid_tmp, 'density_grid_tmp', torch.float32)
    _lib.call('arcb200_splat_grid_samples', density.data_ptr(), density_grid_indices.data_ptr(), int(n_samples),
              density_grid_tmp.data_ptr(), _lib.current_stream())
    return density_grid_tmp


@torch.no_grad()
def ema_grid_samples_nerf(density_grid_tmp, density_grid, n_elements, decay):
    """density_grid = density_grid < 0 ? density_grid : max(density_grid*decay, tmp)   (bitfield_func.py:112-121)"""
    check(density_grid_tmp, 'density_grid_tmp', torch.float32)
    check(density_grid, 'density_grid', torch.float32)
    _lib.call('arcb200_ema_grid_samples_nerf', density_grid_tmp.data_ptr(), int(n_elements), float(decay),
              density_grid.data_ptr(), _lib.current_stream())
    return density_grid


@torch.no_grad()
def update_bitfield(density_grid, density_grid_mean, density_grid_bitfield, thres, n_grid):
    """bit = density > min(thres, mean)   (bitfield_func.py:127-139).  density_grid_mean may be a python
    float (as in the reference) or a 0-dim/1-element CUDA tensor, which avoids the host round trip."""
    check(density_grid, 'density_grid', torch.float32)
    check(density_grid_bitfield, 'density_grid_bitfield', torch.uint8)
    mean_dev = None
    mean_host = 0.0
    if isinstance(density_grid_mean, torch.Tensor) and density_grid_mean.is_cuda:
        mean_dev = f32c(density_grid_mean.reshape(-1)[:1])
    else:
        mean_host = float(density_grid_mean)
    _lib.call('arcb200_update_bitfield', density_grid.data_ptr(), mean_host, _lib.ptr(mean_dev),
              density_grid_bitfield.data_ptr(), float(thres), int(n_grid) ** 3, _lib.current_stream())
    return density_grid_bitfield


@torch.no_grad()
def count_bitfield(density_grid_bitfield, n_grid, faithful=True):
    """Occupied-cell count as a python float (bitfield_func.py:145-159).  faithful=True keeps the
    reference's arithmetic (bitfield_func_kernel.cu:365 uses `&&`, so each non-zero byte counts 8);
    faithful=False returns the true number of set bits."""
    check(density_grid_bitfield, 'density_grid_bitfield', torch.uint8)
    counter = torch.zeros((1, ), dtype=torch.float32, device=density_grid_bitfield.device)
    _lib.call('arcb200_count_bitfield', density_grid_bitfield.data_ptr(), counter.data_ptr(), int(n_grid) ** 3,
              int(bool(faithful)), _lib.current_stream())
    return float(counter[0].item())
