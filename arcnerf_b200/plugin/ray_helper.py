# -*- coding: utf-8 -*-
"""Drop-in for the integration functions of arcnerf/render/ray_helper.py (ray_marching :476-593, alpha_to_weights
:596-620, sample_pdf / sample_cdf :410-473) and arcnerf/models/neus_model.py:sdf_to_alpha (:242-265), differentiable,
on libarcb200's segmented-scan kernels instead of ~15 eager torch ops over padded tensors."""
import os

import torch

from .. import _lib
from .. import functional as F


class _CompositeFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, geo, rad, zvals, noise, bkg, is_alpha, add_inf_z, white_bkg):
        geo_c = geo.contiguous().float()
        rad_c = rad.contiguous().float() if rad is not None else None
        z_c = zvals.contiguous().float()
        if is_alpha and z_c.shape[1] != geo_c.shape[1]:
            raise RuntimeError('alpha and zvals must have the same number of columns')
        bkg_c = bkg.contiguous().float() if bkg is not None else None
        r = F.composite_forward(geo_c, rad_c, z_c, is_alpha=is_alpha, noise=noise, add_inf_z=add_inf_z,
                                white_bkg=white_bkg, bkg=bkg_c, want_weights=True, want_trans=True, want_alpha=True)
        ctx.save_for_backward(geo_c, rad_c if rad_c is not None else torch.empty(0), z_c,
                              noise if noise is not None else torch.empty(0), bkg_c if bkg_c is not None else torch.empty(0))
        ctx.flags = (is_alpha, add_inf_z, white_bkg, rad is not None, noise is not None, bkg is not None)
        rgb = r['rgb'] if r['rgb'] is not None else torch.empty(0, device=geo.device)
        return rgb, r['depth'], r['mask'], r['weights'], r['trans_shift'], r['alpha']

    @staticmethod
    def backward(ctx, g_rgb, g_depth, g_mask, g_w, g_trans, g_alpha):
        geo, rad, z, noise, bkg = ctx.saved_tensors
        is_alpha, add_inf_z, white_bkg, has_rad, has_noise, has_bkg = ctx.flags
        c = lambda g: g.contiguous().float() if g is not None else None
        d_geo, d_rad = F.composite_backward(geo, rad if has_rad else None, z, g_rgb=c(g_rgb) if has_rad else None,
                                            g_depth=c(g_depth), g_mask=c(g_mask), g_weights=c(g_w), is_alpha=is_alpha,
                                            noise=noise if has_noise else None, add_inf_z=add_inf_z, white_bkg=white_bkg,
                                            bkg=bkg if has_bkg else None)
        # gradients through the returned trans_shift / alpha tensors are not propagated (the reference's callers use
        # them for progress records and sampling only, always detached: base_3d_model.py:160-188)
        return d_geo, d_rad, None, None, None, None, None, None


class _DeferredZvalsAssert:
    """ray_marching's `assert torch.all(deltas >= 0), 'zvals is not all increase....'` (ray_helper.py:532-534) without the host
    synchronisation it costs in the reference: the check runs as one small kernel that raises a device flag, the flag travels to
    pinned host memory behind it, and the AssertionError is raised by the first later call that finds the copy complete (or by
    check_zvals_assert(), which waits).  ARCB200_STRICT_ASSERTS=1 waits at once, like the reference."""

    def __init__(self):
        self.state = {}

    def _slot(self, dev):
        if dev not in self.state:
            self.state[dev] = dict(flag=torch.zeros(1, dtype=torch.int32, device=dev), host=torch.zeros(1, dtype=torch.int32).pin_memory(),
                                   event=torch.cuda.Event(), pending=False)
        return self.state[dev]

    def poll(self, dev, wait=False):
        st = self.state.get(dev)
        if st is None or not st['pending']:
            return
        if wait:
            st['event'].synchronize()
        if st['event'].query():
            st['pending'] = False
            if int(st['host'][0]) != 0:
                st['flag'].zero_()
                st['host'].zero_()
                raise AssertionError('zvals is not all increase....')

    def submit(self, zvals):
        if not zvals.is_cuda or zvals.dim() != 2:
            return
        dev = zvals.device
        self.poll(dev)
        st = self._slot(dev)
        z = _f32(zvals.detach())
        with torch.cuda.device(dev):
            _lib.call('arcb200_check_increasing', z.data_ptr(), z.shape[0], z.shape[1], st['flag'].data_ptr(), _lib.current_stream())
            st['host'].copy_(st['flag'], non_blocking=True)
            st['event'].record()
        st['pending'] = True
        if os.environ.get('ARCB200_STRICT_ASSERTS', '0') == '1':
            self.poll(dev, wait=True)


_zvals_assert = _DeferredZvalsAssert()


def check_zvals_assert(device=None):
    """wait for the pending monotonicity checks of ray_marching and raise the reference's AssertionError if one failed"""
    for dev in ([torch.device(device)] if device is not None else list(_zvals_assert.state)):
        _zvals_assert.poll(dev, wait=True)


def ray_marching(sigma, radiance, zvals, add_inf_z=False, noise_std=0.0, weights_only=False, white_bkg=False, alpha=None,
                 bkg_color=None):
    """Same arguments and output dict as arcnerf.render.ray_helper.ray_marching (:476-593)."""
    assert sigma is not None or alpha is not None, 'Can not be None for both alpha and sigma..'
    is_alpha = alpha is not None
    geo = alpha if is_alpha else sigma
    p = zvals.shape[1]
    m = p if (is_alpha or add_inf_z) else p - 1
    _zvals_assert.submit(zvals)
    noise = None
    if not is_alpha and noise_std > 0.0:
        noise = torch.randn(geo.shape, dtype=torch.float32, device=geo.device) * noise_std
    rgb, depth, mask, weights, trans, alpha_out = _CompositeFn.apply(geo, radiance, zvals, noise, bkg_color, is_alpha,
                                                                     bool(add_inf_z), bool(white_bkg))
    weights = weights[:, :m]
    if weights_only:
        return {'weights': weights}
    _sigma, _radiance, _zvals = sigma, radiance, zvals
    if not add_inf_z and not is_alpha:
        _sigma = sigma[:, :-1] if sigma is not None else None
        _radiance = radiance[:, :-1, :] if radiance is not None else None
        _zvals = zvals[:, :-1]
    return {
        'rgb': rgb if radiance is not None else None, 'depth': depth, 'mask': mask, 'sigma': _sigma,
        'radiance': _radiance, 'zvals': _zvals, 'alpha': alpha if is_alpha else alpha_out[:, :m],
        'trans_shift': trans[:, :m], 'weights': weights,
    }


def alpha_to_weights(alpha):
    """-> trans_shift, weights (N_rays, N_p)   (ray_helper.py:596-620)"""
    z = torch.zeros_like(alpha)
    _, _, _, weights, trans, _ = _CompositeFn.apply(alpha, None, z, None, None, True, False, False)
    return trans, weights


def sample_cdf(bins, cdf, n_sample, det=False, eps=1e-5):
    """ray_helper.py:434-473: uniform (det) or random u, inverse CDF by searchsorted + interpolation, sorted."""
    if det:
        u = torch.linspace(0.0, 1.0, steps=n_sample, device=bins.device)
    else:
        u = torch.rand(list(cdf.shape[:-1]) + [n_sample], device=bins.device)
    with torch.no_grad():
        samples = F.sample_cdf_gather(bins.contiguous().float(), cdf.detach().contiguous().float(), u.contiguous(), eps)
        samples = sort_rows(samples)
    return samples


def sample_pdf(bins, weights, n_sample, det=False, eps=1e-5):
    """ray_helper.py:410-431"""
    weights = weights + eps
    pdf = weights / torch.sum(weights, -1, keepdim=True)
    cdf = torch.cumsum(pdf, -1)
    cdf = torch.cat([torch.zeros_like(cdf[..., :1]), cdf], -1)
    return sample_cdf(bins, cdf, n_sample, det)


class _SdfToAlphaFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, sdf, zvals, slope, s, clip):
        sdf_c, z_c, sl_c = sdf.contiguous().float(), zvals.contiguous().float(), slope.contiguous().float()
        s_c = s.detach().reshape(1).float().contiguous() if isinstance(s, torch.Tensor) else float(s)
        ctx.save_for_backward(sdf_c, z_c, sl_c, s_c if isinstance(s_c, torch.Tensor) else torch.empty(0))
        ctx.s_host, ctx.clip, ctx.s_shape = (None if isinstance(s_c, torch.Tensor) else s_c), clip, \
            (s.shape if isinstance(s, torch.Tensor) else None)
        return F.sdf_to_alpha(sdf_c, z_c, sl_c, s_c, clip)

    @staticmethod
    def backward(ctx, g):
        sdf, z, slope, s = ctx.saved_tensors
        sv = s if ctx.s_host is None else ctx.s_host
        d_sdf, d_slope, d_s = F.sdf_to_alpha_backward(sdf, z, slope, sv, g.contiguous().float(), ctx.clip)
        return d_sdf, None, d_slope, (d_s.reshape(ctx.s_shape) if ctx.s_shape is not None else None), None


def sdf_to_alpha(mid_sdf, zvals, mid_slope, s, clip=True):
    """arcnerf/models/neus_model.py:242-265; differentiable w.r.t. mid_sdf, mid_slope and the scale s."""
    return _SdfToAlphaFn.apply(mid_sdf, zvals, mid_slope, s, clip)


# ------------------------------------------------------------------------------------------ depth samples (a15)
def _f32(t):
    return t.contiguous() if t.dtype == torch.float32 else t.float().contiguous()


def sort_rows(vals):
    """ascending sort of every row (torch.sort(vals, -1)[0] of the reference's sample merge)"""
    v = _f32(vals)
    flat = v.reshape(-1, v.shape[-1])
    out = torch.empty_like(flat)
    _lib.call('arcb200_sort_rows', flat.data_ptr(), out.data_ptr(), flat.shape[0], flat.shape[1], _lib.current_stream())
    return out.reshape(v.shape)


@torch.no_grad()
def sphere_ray_intersection(rays_o, rays_d, radius, origin=(0, 0, 0)):
    """arcnerf/geometry/ray.py:180-255 -> near, far (N_rays, N_r), pts (N_rays, N_r, 2, 3), mask (N_rays, N_r) bool"""
    rays_o, rays_d = _f32(rays_o), _f32(rays_d)
    if not isinstance(radius, torch.Tensor):
        assert isinstance(radius, (float, int)), 'Invalid type'
        radius = torch.tensor([radius], dtype=torch.float32, device=rays_o.device)
    radius = _f32(radius).to(rays_o.device)
    n, nr = rays_o.shape[0], radius.shape[0]
    dev = rays_o.device
    near = torch.empty((n, nr), dtype=torch.float32, device=dev)
    far = torch.empty((n, nr), dtype=torch.float32, device=dev)
    pts = torch.empty((n, nr, 2, 3), dtype=torch.float32, device=dev)
    mask = torch.empty((n, nr), dtype=torch.bool, device=dev)
    import ctypes
    org = (ctypes.c_float * 3)(*[float(x) for x in origin])
    _lib.call('arcb200_sphere_ray_intersection', rays_o.data_ptr(), rays_d.data_ptr(), radius.data_ptr(), org, near.data_ptr(),
              far.data_ptr(), pts.data_ptr(), mask.data_ptr(), n, nr, _lib.current_stream())
    return near, far, pts, mask


@torch.no_grad()
def get_ray_points_by_zvals(rays_o, rays_d, zvals):
    """arcnerf/geometry/ray.py:11-30 -> (N_rays, N_pts, 3)"""
    rays_o, rays_d, zvals = _f32(rays_o), _f32(rays_d), _f32(zvals)
    assert zvals.shape[0] == rays_o.shape[0], 'Invalid shape for zvals... Should be (N_rays, N_pts)'
    pts = torch.empty((zvals.shape[0], zvals.shape[1], 3), dtype=torch.float32, device=zvals.device)
    _lib.call('arcb200_ray_points', rays_o.data_ptr(), rays_d.data_ptr(), zvals.data_ptr(), pts.data_ptr(), zvals.shape[0],
              zvals.shape[1], _lib.current_stream())
    return pts


@torch.no_grad()
def get_near_far_from_rays(rays_o, rays_d, bounds=None, near_hardcode=None, far_hardcode=None, bounding_radius=None):
    """ray_helper.py:175-228 -> near, far (N_rays, 1)"""
    dev, n_rays = rays_o.device, rays_o.shape[0]
    if near_hardcode is None or far_hardcode is None:
        if bounds is None and bounding_radius is None:
            raise NotImplementedError('You must specify near/far in some place...')
        if bounds is None:
            near, far, _, _ = sphere_ray_intersection(rays_o, rays_d, float(bounding_radius))
        else:
            near, far = bounds[:, 0:1].float().clone(), bounds[:, 1:2].float().clone()
            if bounding_radius is not None:
                _, far_bound, _, _ = sphere_ray_intersection(rays_o, rays_d, float(bounding_radius))
                far = torch.where(far > far_bound, far_bound, far)
        if near_hardcode is not None:
            near = near * 0 + near_hardcode
        if far_hardcode is not None:
            far = far * 0 + far_hardcode
    else:
        near = torch.full((n_rays, 1), float(near_hardcode), dtype=torch.float32, device=dev)
        far = torch.full((n_rays, 1), float(far_hardcode), dtype=torch.float32, device=dev)
    far = torch.where(far <= near, near + 1e-5, far)
    return near, far


@torch.no_grad()
def get_zvals_from_near_far(near, far, n_pts, inclusive=True, inverse_linear=False, perturb=False, u=None):
    """ray_helper.py:231-263 -> (N_rays, n_pts).  u: optional uniforms (N_rays, n_pts) for the jitter (default torch.rand)."""
    near, far = _f32(near).reshape(-1), _f32(far).reshape(-1)
    n = near.shape[0]
    if perturb and u is None:
        u = torch.rand((n, n_pts), dtype=torch.float32, device=near.device)
    zvals = torch.empty((n, n_pts), dtype=torch.float32, device=near.device)
    _lib.call('arcb200_zvals_from_near_far', near.data_ptr(), far.data_ptr(), int(n_pts), int(bool(inclusive)),
              int(bool(inverse_linear)), _lib.ptr(_f32(u) if (perturb and u is not None) else None), zvals.data_ptr(), n,
              _lib.current_stream())
    return zvals


@torch.no_grad()
def get_zvals_from_near_far_fix_step(near, far, fix_t, n_pts, inclusive=True, perturb=False, u=None):
    """ray_helper.py:266-315 -> zvals (N_rays, n_pts), mask_pts (N_rays, n_pts) bool; jitters always, as the reference"""
    assert fix_t > 0, 'Only allow positive step...'
    near, far = _f32(near).reshape(-1), _f32(far).reshape(-1)
    n = near.shape[0]
    if u is None:
        u = torch.rand((n, n_pts), dtype=torch.float32, device=near.device)
    zvals = torch.empty((n, n_pts), dtype=torch.float32, device=near.device)
    mask = torch.empty((n, n_pts), dtype=torch.bool, device=near.device)
    _lib.call('arcb200_zvals_fix_step', near.data_ptr(), far.data_ptr(), float(fix_t), int(n_pts), int(bool(inclusive)),
              _f32(u).data_ptr(), zvals.data_ptr(), mask.data_ptr(), n, _lib.current_stream())
    return zvals, mask


@torch.no_grad()
def perturb_interval(vals, u=None):
    """ray_helper.py:361-378 on any (B, N) values (kept in torch: four elementwise ops on a small tensor)"""
    mids = .5 * (vals[..., 1:] + vals[..., :-1])
    upper = torch.cat([mids, vals[..., -1:]], -1)
    lower = torch.cat([vals[..., :1], mids], -1)
    if u is None:
        u = torch.rand(upper.shape, dtype=vals.dtype, device=vals.device)
    return lower + (upper - lower) * u


@torch.no_grad()
def get_zvals_from_sphere_radius(rays_o, rays_d, sphere_radius):
    """ray_helper.py:344-358: far intersection with every shell -> (N_rays, N)"""
    _, zvals, _, _ = sphere_ray_intersection(rays_o, rays_d, sphere_radius)
    return zvals


@torch.no_grad()
def get_zvals_outside_sphere(rays_o, rays_d, n_pts, radius, perturb=False, u=None):
    """ray_helper.py:318-341 (NeRF++ inverse-sphere shells) -> zvals (N_rays, n_pts), sphere_radius (n_pts,)"""
    t_vals = torch.linspace(0.0, 1.0, n_pts + 2, dtype=torch.float32, device=rays_o.device)[1:-1]
    sphere_radius = radius / torch.flip(t_vals, dims=[-1])
    if perturb:
        sphere_radius = perturb_interval(sphere_radius[None], u)[0]
    return get_zvals_from_sphere_radius(rays_o, rays_d, sphere_radius), sphere_radius


@torch.no_grad()
def upsample_zvals(zvals, weights, n_importance, det=True, eps=1e-5):
    """NeRF.upsample_zvals (arcnerf/models/nerf_model.py:93-117) without the mask bookkeeping: importance samples from
    the coarse weights on the interval mid-points, merged into the coarse samples in ascending order."""
    n_sample = zvals.shape[1]
    weights_coarse = weights[:, 1:n_sample - 1]
    zvals_mid = 0.5 * (zvals[..., 1:] + zvals[..., :-1])
    new = sample_pdf(zvals_mid, weights_coarse, n_importance, det, eps).detach()
    return sort_rows(torch.cat([zvals, new], -1)), new


def pack_cameras(intrinsics, c2ws, device):
    """(n,3,3) intrinsics + (n,4,4) poses -> the (n,17) device table arcb200_gen_rays reads"""
    k, c = torch.as_tensor(intrinsics, dtype=torch.float32).reshape(-1, 3, 3), torch.as_tensor(c2ws, dtype=torch.float32).reshape(-1, 4, 4)
    tab = torch.cat([k[:, 0, 0:1], k[:, 1, 1:2], k[:, 0, 2:3], k[:, 1, 2:3], k[:, 0, 1:2], c[:, :3, :].reshape(-1, 12)], 1)
    return tab.contiguous().to(device)


@torch.no_grad()
def get_rays(W, H, intrinsic, c2w, wh_order=True, index=None, n_rays=-1, to_np=False, ndc=False, ndc_near=1.0, center_pixel=False,
             normalize_rays_d=True):
    """arcnerf.render.ray_helper.get_rays (:12-119) on the device: same arguments and returns (rays_o, rays_d, index list,
    rays_r).  NDC rays are not provided."""
    assert (index is None) or n_rays <= 0, 'You are not allowed to sampled both by index and N_ray'
    assert not ndc, 'ndc rays are not provided by this build'
    dev = intrinsic.device
    if not intrinsic.is_cuda:
        raise RuntimeError('get_rays needs CUDA tensors')
    cams = pack_cameras(intrinsic, c2w, dev)
    pix = None
    if index is not None:
        assert len(index.shape) == 2 and index.shape[-1] == 2, 'invalid shape, should be (N_rays, 2)'
        idx = torch.as_tensor(index, dtype=torch.long, device=dev)
        pix = (idx[:, 0] * H + idx[:, 1]).contiguous()
    if n_rays > 0:
        pix = torch.randperm(W * H, device=dev)[:n_rays].contiguous()
    n = W * H if pix is None else pix.shape[0]
    rays_o = torch.empty((n, 3), dtype=torch.float32, device=dev)
    rays_d = torch.empty((n, 3), dtype=torch.float32, device=dev)
    rays_r = torch.empty((n, 1), dtype=torch.float32, device=dev) if pix is None else None
    _lib.call('arcb200_gen_rays', cams.data_ptr(), 1, None, _lib.ptr(pix), int(W), int(H), int(bool(wh_order)), int(bool(center_pixel)),
              int(bool(normalize_rays_d)), rays_o.data_ptr(), rays_d.data_ptr(), _lib.ptr(rays_r), n, _lib.current_stream())
    index_list = pix.tolist() if pix is not None else None
    if to_np:
        return rays_o.cpu().numpy(), rays_d.cpu().numpy(), index_list, rays_r
    return rays_o, rays_d, index_list, rays_r


@torch.no_grad()
def get_train_batch_rays(cams, cam_id, pix, W, H, center_pixel=False, normalize_rays_d=True):
    """A training batch drawn from many cameras in one launch (what Pipeline.get_train_batch gathers from pre-computed per-image
    rays, pipeline.py:204-300): cams from pack_cameras, cam_id (N) int32, pix (N) int64 pixel index i * H + j."""
    n = pix.shape[0]
    rays_o = torch.empty((n, 3), dtype=torch.float32, device=pix.device)
    rays_d = torch.empty((n, 3), dtype=torch.float32, device=pix.device)
    _lib.call('arcb200_gen_rays', cams.data_ptr(), cams.shape[0], cam_id.data_ptr(), pix.data_ptr(), int(W), int(H), 1,
              int(bool(center_pixel)), int(bool(normalize_rays_d)), rays_o.data_ptr(), rays_d.data_ptr(), None, n, _lib.current_stream())
    return rays_o, rays_d
