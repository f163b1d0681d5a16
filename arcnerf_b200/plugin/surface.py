# -*- coding: utf-8 -*-
"""Surface-ray intersection: arcnerf/geometry/ray.py:352-601 surface_ray_intersection / sphere_tracing /
secant_root_finding with the same arguments and returns, on the library's per-ray kernels (csrc/surface.cu).  `geo_func` is the
caller's field: (N_pts, 3) CUDA tensor -> (N_pts,) values (for the SDF models: GeoNet.forward_geo_value, which runs on the
library's hash-grid / dense kernels).  No host synchronisation happens inside the loops."""
import torch

from .. import _lib


def _near_far(v, n_rays, dev):
    """(n_rays, 1) tensors are used per ray, anything else is one value for all rays (ray.py:429-436)"""
    if isinstance(v, torch.Tensor) and tuple(v.shape) == (n_rays, 1):
        return v.to(dev, torch.float32).contiguous().view(-1), 1
    return torch.full((1, ), float(v), dtype=torch.float32, device=dev), 0


def _check(rays_o, rays_d):
    if not (rays_o.is_cuda and rays_d.is_cuda):
        raise RuntimeError('rays must be CUDA tensors')
    return rays_o.contiguous().float(), rays_d.contiguous().float()


@torch.no_grad()
def sphere_tracing(rays_o, rays_d, sdf_func, near=0.0, far=10.0, n_iter=100, threshold=0.001):
    """ray.py:403-467 -> zvals (N_rays, 1), pts (N_rays, 3), mask (N_rays,) bool"""
    rays_o, rays_d = _check(rays_o, rays_d)
    n, dev = rays_o.shape[0], rays_o.device
    nr, per_n = _near_far(near, n, dev)
    fr, per_f = _near_far(far, n, dev)
    if per_n != per_f:                                     # mixed scalar / per-ray bounds: expand the scalar one
        nr = nr.expand(n).contiguous() if not per_n else nr
        fr = fr.expand(n).contiguous() if not per_f else fr
    per = int(per_n or per_f)
    zvals = (nr if per else nr.expand(n)).clone().contiguous()
    sdf = torch.zeros(n, dtype=torch.float32, device=dev)
    mask = torch.ones(n, dtype=torch.uint8, device=dev)
    obj = torch.zeros(n, dtype=torch.uint8, device=dev)
    pts = torch.empty((n, 3), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        st = _lib.current_stream()
        for _ in range(int(n_iter)):
            _lib.call('arcb200_ray_points', rays_o.data_ptr(), rays_d.data_ptr(), zvals.data_ptr(), pts.data_ptr(), n, 1, st)
            val = sdf_func(pts).reshape(-1).float().contiguous()
            _lib.call('arcb200_trace_update', val.data_ptr(), zvals.data_ptr(), sdf.data_ptr(), nr.data_ptr(), fr.data_ptr(), per,
                      mask.data_ptr(), obj.data_ptr(), float(threshold), n, st)
        _lib.call('arcb200_surface_finish', rays_o.data_ptr(), rays_d.data_ptr(), zvals.data_ptr(), nr.data_ptr(), per, pts.data_ptr(), n, st)
    return zvals.view(n, 1), pts, mask.bool()


@torch.no_grad()
def secant_root_finding(rays_o, rays_d, geo_func, near=0.0, far=10.0, n_step=128, n_iter=20, threshold=0.001, level=0.0,
                        grad_dir='ascent'):
    """ray.py:470-601 -> zvals (N_rays, 1), pts (N_rays, 3), mask (N_rays,) bool"""
    rays_o, rays_d = _check(rays_o, rays_d)
    n, dev = rays_o.shape[0], rays_o.device
    nr, per_n = _near_far(near, n, dev)
    fr, per_f = _near_far(far, n, dev)
    if per_n != per_f:
        nr = nr.expand(n).contiguous() if not per_n else nr
        fr = fr.expand(n).contiguous() if not per_f else fr
    per, descent = int(per_n or per_f), int(grad_dir == 'descent')
    f32, u8 = torch.float32, torch.uint8
    e = lambda *shape, dt=f32: torch.empty(shape, dtype=dt, device=dev)
    step, scan_pts = e(n, n_step), e(n * n_step, 3)
    not_occ, mask = e(n, dt=u8), e(n, dt=u8)
    z_high, z_low, g_high, g_low, z_mid, z_init, zvals = (e(n) for _ in range(7))
    flags = torch.empty(int(n_iter) + 1, dtype=torch.int32, device=dev)
    pts = e(n, 3)
    with torch.cuda.device(dev):
        st = _lib.current_stream()
        _lib.call('arcb200_secant_steps', rays_o.data_ptr(), rays_d.data_ptr(), nr.data_ptr(), fr.data_ptr(), per, int(n_step),
                  step.data_ptr(), scan_pts.data_ptr(), n, st)
        geo = geo_func(scan_pts).reshape(-1).float().contiguous()
        _lib.call('arcb200_secant_init', geo.data_ptr(), step.data_ptr(), int(n_step), float(level), descent, not_occ.data_ptr(),
                  mask.data_ptr(), z_high.data_ptr(), z_low.data_ptr(), g_high.data_ptr(), g_low.data_ptr(), z_mid.data_ptr(),
                  z_init.data_ptr(), flags.data_ptr(), int(n_iter), n, st)
        for it in range(int(n_iter)):
            _lib.call('arcb200_ray_points', rays_o.data_ptr(), rays_d.data_ptr(), z_mid.data_ptr(), pts.data_ptr(), n, 1, st)
            val = geo_func(pts).reshape(-1).float().contiguous()
            _lib.call('arcb200_secant_update', val.data_ptr(), float(level), descent, mask.data_ptr(), z_low.data_ptr(), g_low.data_ptr(),
                      z_high.data_ptr(), g_high.data_ptr(), z_mid.data_ptr(), z_init.data_ptr(), float(threshold), flags.data_ptr(), it,
                      int(n_iter), n, st)
        _lib.call('arcb200_secant_select', mask.data_ptr(), not_occ.data_ptr(), z_mid.data_ptr(), fr.data_ptr(), per, zvals.data_ptr(), n, st)
        _lib.call('arcb200_surface_finish', rays_o.data_ptr(), rays_d.data_ptr(), zvals.data_ptr(), nr.data_ptr(), per, pts.data_ptr(), n, st)
    return zvals.view(n, 1), pts, mask.bool()


def surface_ray_intersection(rays_o, rays_d, geo_func, method='sphere_tracing', near=0.0, far=10.0, n_step=128, n_iter=100,
                             threshold=0.001, level=0.0, grad_dir='ascent'):
    """ray.py:352-400"""
    if method == 'sphere_tracing':
        return sphere_tracing(rays_o, rays_d, geo_func, near, far, n_iter, threshold)
    if method == 'secant_root_finding':
        return secant_root_finding(rays_o, rays_d, geo_func, near, far, n_step, n_iter, threshold, level, grad_dir)
    raise NotImplementedError('Method {} not support for surface-ray intersection'.format(method))
