# -*- coding: utf-8 -*-
"""Drop-in for arcnerf/ops/multivol_func.py (cascaded multi-volume sampler used by the MultiVol
background model), backed by libarcb200 instead of `_multivol_func` (multivol_func.py:17-143)."""
import torch

from .. import _lib
from ._common import check, f32c
from .volume_func import _sampling_args

CUDA_BACKEND_AVAILABLE = True

rng = None


def _rng():
    global rng
    if rng is None:
        rng = _lib.Rng()
    return rng


def reset_rng(seed=_lib.Rng.SEED):
    global rng
    rng = _lib.Rng(seed)


@torch.no_grad()
def sparse_sampling_in_multivol_bitfield(
    rays_o, rays_d, near, far, n_pts, cone_angle, min_step, max_step, min_aabb_range, aabb_range, n_grid, n_cascade,
    bitfield, near_distance=0.0, inclusive=False
):
    """Cone-step march through n_cascade nested volumes (multivol_func.py:47-89) -> zvals, mask (N,n_pts)."""
    rays_o, rays_d, near, far, aabb, n_rays = _sampling_args(rays_o, rays_d, near, far, aabb_range)
    min_aabb = check(f32c(torch.permute(min_aabb_range, (1, 0))), 'min_aabb_range', torch.float32, (2, 3))
    check(bitfield, 'bitfield', torch.uint8)
    n_level = n_cascade if inclusive else n_cascade - 1
    if bitfield.numel() != n_grid ** 3 // 8 * n_level:
        raise RuntimeError('bitfield must hold n_grid**3 / 8 bytes per stored cascade')
    zvals = torch.empty((n_rays, n_pts), dtype=torch.float32, device=rays_o.device)
    mask = torch.empty((n_rays, n_pts), dtype=torch.bool, device=rays_o.device)
    r = _rng()
    _lib.call('arcb200_sparse_sampling_in_multivol_bitfield', rays_o.data_ptr(), rays_d.data_ptr(), near.data_ptr(),
              far.data_ptr(), int(n_pts), float(cone_angle), float(min_step), float(max_step), min_aabb.data_ptr(),
              aabb.data_ptr(), int(n_grid), int(n_cascade), bitfield.data_ptr(), float(near_distance),
              int(bool(inclusive)), zvals.data_ptr(), mask.data_ptr(), n_rays, r.state, r.inc, _lib.current_stream())
    r.advance()
    return zvals, mask


@torch.no_grad()
def generate_grid_samples_multivol(
    density_grid, n_elements, aabb_range, density_grid_ema_step, n_cascade, n_grid, thresh, inclusive
):
    """-> world positions (n_elements,3), indices (n_elements,) int32   (multivol_func.py:95-125)"""
    check(density_grid, 'density_grid', torch.float32)
    aabb = check(f32c(torch.permute(aabb_range, (1, 0))), 'aabb_range', torch.float32, (2, 3))
    dev = density_grid.device
    positions = torch.empty((n_elements, 3), dtype=torch.float32, device=dev)
    indices = torch.empty((n_elements, ), dtype=torch.int32, device=dev)
    r = _rng()
    _lib.call('arcb200_generate_grid_samples_multivol', density_grid.data_ptr(), int(density_grid_ema_step),
              int(n_elements), aabb.data_ptr(), int(n_cascade), int(n_grid), float(thresh), int(bool(inclusive)),
              positions.data_ptr(), indices.data_ptr(), r.state, r.inc, _lib.current_stream())
    r.advance()
    return positions, indices


@torch.no_grad()
def update_bitfield_multivol(density_grid, density_grid_mean, density_grid_bitfield, thres, n_grid, n_cascade, inclusive):
    """bit = density > min(thres, mean) over all stored cascades   (multivol_func.py:131-143)"""
    check(density_grid, 'density_grid', torch.float32)
    check(density_grid_bitfield, 'density_grid_bitfield', torch.uint8)
    n_level = n_cascade if inclusive else n_cascade - 1
    _lib.call('arcb200_update_bitfield', density_grid.data_ptr(), float(density_grid_mean), None,
              density_grid_bitfield.data_ptr(), float(thres), int(n_grid) ** 3 * n_level, _lib.current_stream())
    return density_grid_bitfield
