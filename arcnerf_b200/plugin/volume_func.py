# -*- coding: utf-8 -*-
"""Drop-in for arcnerf/ops/volume_func.py, backed by libarcb200 instead of the `_volume_func` pybind module.

Same function names, arguments and returns (arcnerf/ops/volume_func.py:15-167).  Differences, all
invisible to callers: outputs are allocated with torch.empty (the kernels write every entry), work is
enqueued on torch's current stream and nothing synchronises the device (the reference launches on the
legacy stream and calls cudaDeviceSynchronize, volume_func_kernel.cu:55,67).
"""
import torch

from .. import _lib
from ._common import check, f32c

CUDA_BACKEND_AVAILABLE = True

# The reference keeps one `static pcg32 rng{9121}` per extension module (include/common.h:21-23).
rng = None


def _rng():
    global rng
    if rng is None:
        rng = _lib.Rng()
    return rng


def reset_rng(seed=_lib.Rng.SEED):
    """Re-seed this module's generator (a fresh import of the reference module starts from 9121)."""
    global rng
    rng = _lib.Rng(seed)


@torch.no_grad()
def check_pts_in_occ_voxel_cuda(xyz, bitfield, aabb_range, n_grid):
    """xyz (B,3); bitfield (n,n,n) bool; aabb_range (3,2) -> (B,) bool   (volume_func.py:30-43)"""
    xyz = check(f32c(xyz), 'xyz', torch.float32, (None, 3))
    check(bitfield, 'bitfield', torch.bool)
    if bitfield.numel() != n_grid ** 3:
        raise RuntimeError('bitfield must hold n_grid**3 entries')
    aabb = check(f32c(torch.permute(aabb_range, (1, 0))), 'aabb_range', torch.float32, (2, 3))
    out = torch.empty((xyz.shape[0], ), dtype=torch.bool, device=xyz.device)
    _lib.call('arcb200_check_pts_in_occ_voxel', xyz.data_ptr(), bitfield.data_ptr(), aabb.data_ptr(), int(n_grid),
              out.data_ptr(), xyz.shape[0], _lib.current_stream())
    return out


@torch.no_grad()
def ray_aabb_intersection_cuda(rays_o, rays_d, aabb_range):
    """rays (N,3); aabb_range (N_v,3,2) -> near, far (N,N_v), pts (N,N_v,2,3), mask (N,N_v) bool
    (volume_func.py:68-84).  A ray whose origin lies inside a box misses it, as in the reference kernel."""
    rays_o = check(f32c(rays_o), 'rays_o', torch.float32, (None, 3))
    rays_d = check(f32c(rays_d), 'rays_d', torch.float32, (rays_o.shape[0], 3))
    aabb = check(f32c(torch.permute(aabb_range, (0, 2, 1))), 'aabb_range', torch.float32, (None, 2, 3))
    n_rays, n_v = rays_o.shape[0], aabb.shape[0]
    dev = rays_o.device
    near = torch.empty((n_rays, n_v), dtype=torch.float32, device=dev)
    far = torch.empty((n_rays, n_v), dtype=torch.float32, device=dev)
    pts = torch.empty((n_rays, n_v, 2, 3), dtype=torch.float32, device=dev)
    mask = torch.empty((n_rays, n_v), dtype=torch.bool, device=dev)
    _lib.call('arcb200_aabb_intersection', rays_o.data_ptr(), rays_d.data_ptr(), aabb.data_ptr(), near.data_ptr(),
              far.data_ptr(), pts.data_ptr(), mask.data_ptr(), n_rays, n_v, _lib.current_stream())
    return near, far, pts, mask


def _sampling_args(rays_o, rays_d, near, far, aabb_range):
    rays_o = check(f32c(rays_o), 'rays_o', torch.float32, (None, 3))
    n_rays = rays_o.shape[0]
    rays_d = check(f32c(rays_d), 'rays_d', torch.float32, (n_rays, 3))
    near = check(f32c(near), 'near', torch.float32)
    far = check(f32c(far), 'far', torch.float32)
    if near.numel() != n_rays or far.numel() != n_rays:
        raise RuntimeError('near/far must be (N_rays, 1)')
    aabb = check(f32c(torch.permute(aabb_range, (1, 0))), 'aabb_range', torch.float32, (2, 3))
    return rays_o, rays_d, near, far, aabb, n_rays


@torch.no_grad()
def sparse_volume_sampling(rays_o, rays_d, near, far, n_pts, dt, aabb_range, n_grid, bitfield, near_distance=0.0):
    """Fixed-step march through a bool occupancy volume -> zvals (N,n_pts), mask (N,n_pts) bool
    (volume_func.py:110-134): valid samples first, then the last valid distance with mask False."""
    rays_o, rays_d, near, far, aabb, n_rays = _sampling_args(rays_o, rays_d, near, far, aabb_range)
    check(bitfield, 'bitfield', torch.bool)
    if bitfield.numel() != n_grid ** 3:
        raise RuntimeError('bitfield must be (n_grid, n_grid, n_grid)')
    zvals = torch.empty((n_rays, n_pts), dtype=torch.float32, device=rays_o.device)
    mask = torch.empty((n_rays, n_pts), dtype=torch.bool, device=rays_o.device)
    r = _rng()
    _lib.call('arcb200_sparse_volume_sampling', rays_o.data_ptr(), rays_d.data_ptr(), near.data_ptr(), far.data_ptr(),
              int(n_pts), float(dt), aabb.data_ptr(), int(n_grid), bitfield.data_ptr(), float(near_distance),
              zvals.data_ptr(), mask.data_ptr(), n_rays, r.state, r.inc, _lib.current_stream())
    r.advance()
    return zvals, mask


@torch.no_grad()
def tensor_reduce_max(full_tensor, idx, n_group):
    """Max of full_tensor (N,) per group idx (N,) int64 -> (n_group,)   (volume_func.py:153-167)"""
    full_tensor = check(f32c(full_tensor), 'full_tensor', torch.float32, (None, ))
    idx = check(idx.contiguous(), 'idx', torch.int64, (full_tensor.shape[0], ))
    uni = torch.zeros((n_group, ), dtype=torch.float32, device=full_tensor.device)
    _lib.call('arcb200_tensor_reduce_max', full_tensor.data_ptr(), idx.data_ptr(), int(n_group), uni.data_ptr(),
              full_tensor.shape[0], _lib.current_stream())
    return uni
