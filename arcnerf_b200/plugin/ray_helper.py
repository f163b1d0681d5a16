# -*- coding: utf-8 -*-
"""Drop-in for the integration functions of arcnerf/render/ray_helper.py (ray_marching :476-593, alpha_to_weights
:596-620, sample_pdf / sample_cdf :410-473) and arcnerf/models/neus_model.py:sdf_to_alpha (:242-265), differentiable,
on libarcb200's segmented-scan kernels instead of ~15 eager torch ops over padded tensors."""
import torch

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


def ray_marching(sigma, radiance, zvals, add_inf_z=False, noise_std=0.0, weights_only=False, white_bkg=False, alpha=None,
                 bkg_color=None):
    """Same arguments and output dict as arcnerf.render.ray_helper.ray_marching (:476-593)."""
    assert sigma is not None or alpha is not None, 'Can not be None for both alpha and sigma..'
    is_alpha = alpha is not None
    geo = alpha if is_alpha else sigma
    p = zvals.shape[1]
    m = p if (is_alpha or add_inf_z) else p - 1
    with torch.no_grad():
        deltas = zvals[:, 1:] - zvals[:, :-1]
        deltas = torch.where(deltas.abs() < 1e-5, torch.zeros_like(deltas), deltas)
        assert torch.all(deltas >= 0), 'zvals is not all increase....'
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
        samples, _ = torch.sort(samples, -1)
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
