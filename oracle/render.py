"""CPU oracle for the ray-integration functions of the hot path.  TEST INFRASTRUCTURE ONLY.

Plain torch-CPU / numpy restatement of arcnerf/render/ray_helper.py (ray_marching :476-593, alpha_to_weights
:596-620, sample_pdf / sample_cdf :410-473) and arcnerf/models/neus_model.py (sdf_to_alpha :242-265).  Pinned
against outputs of the reference's own Python functions, imported in the build container by
tests/golden/make_golden_torch.py and committed as tests/golden/render_golden.npz.
"""
import numpy as np
import torch


def alpha_to_weights(alpha):
    """ray_helper.py:596-620"""
    one = torch.ones_like(alpha[:, :1])
    trans = torch.cat([one, 1 - alpha + 1e-10], -1)
    trans = torch.cumprod(trans, -1)[:, :-1]
    return trans, alpha * trans


def ray_marching(sigma, radiance, zvals, add_inf_z=False, noise=None, white_bkg=False, alpha=None, bkg_color=None):
    """ray_helper.py:476-593 with the random noise passed in explicitly (already scaled by noise_std)."""
    n_rays = zvals.shape[0]
    deltas = zvals[:, 1:] - zvals[:, :-1]
    deltas = torch.where(deltas.abs() < 1e-5, torch.zeros_like(deltas), deltas)
    _sigma, _radiance, _zvals = sigma, radiance, zvals
    if add_inf_z:
        deltas = torch.cat([deltas, 1e10 * torch.ones((n_rays, 1), dtype=zvals.dtype)], -1)
    elif alpha is None:
        _sigma = sigma[:, :-1] if sigma is not None else None
        _radiance = radiance[:, :-1, :] if radiance is not None else None
        _zvals = zvals[:, :-1]
    if alpha is None:
        s = _sigma if noise is None else _sigma + noise[:, :_sigma.shape[1]]
        alpha = 1 - torch.exp(-torch.relu(s) * deltas)
    trans, weights = alpha_to_weights(alpha)
    depth = torch.sum(weights * _zvals, -1)
    mask = torch.sum(weights, -1)
    rgb = None
    if _radiance is not None:
        rgb = torch.sum(weights.unsqueeze(-1) * _radiance, -2)
        if bkg_color is not None:
            rgb = rgb + trans[:, -1:] * bkg_color
        elif white_bkg:
            rgb = rgb + (1.0 - mask[:, None])
    return dict(rgb=rgb, depth=depth, mask=mask, alpha=alpha, trans_shift=trans, weights=weights)


def pad_packed(values, counts, pad_last=True):
    """Packed per-sample array -> the padded (R, P_max) matrix FgModel builds (fg_model.py:311-316): every row is
    filled with the ray's last valid value, then the valid prefix is written.  P_max = max(2, counts.max())
    (reduce_empty_mask, fg_model.py:251-262).  Rows of empty rays are zero."""
    counts = np.asarray(counts)
    r = counts.shape[0]
    p = max(2, int(counts.max()) if r else 2)
    values = np.asarray(values)
    out = np.zeros((r, p) + values.shape[1:], values.dtype)
    off = np.concatenate([[0], np.cumsum(counts)])
    for i in range(r):
        c = counts[i]
        if c:
            out[i, :] = values[off[i] + c - 1]
            out[i, :c] = values[off[i]:off[i] + c]
    return out


def sdf_to_alpha(mid_sdf, zvals, mid_slope, s, clip=True):
    """neus_model.py:242-265"""
    dist = zvals[:, 1:] - zvals[:, :-1]
    prev = torch.sigmoid((mid_sdf - mid_slope * dist * 0.5) * s)
    nxt = torch.sigmoid((mid_sdf + mid_slope * dist * 0.5) * s)
    a = (prev - nxt + 1e-5) / (prev + 1e-5)
    return a.clip(0.0, 1.0) if clip else a


def sample_cdf(bins, cdf, u, eps=1e-5):
    """ray_helper.py:434-473 with the uniform samples u (B, n_sample) passed in; returns SORTED samples."""
    n_pts = bins.shape[-1]
    n_sample = u.shape[-1]
    inds = torch.searchsorted(cdf, u.contiguous(), right=True)
    below = torch.clamp(inds - 1, 0, n_pts - 1)
    above = torch.clamp(inds, 0, n_pts - 1)
    inds_g = torch.stack([below, above], -1).view(-1, 2 * n_sample)
    cdf_g = torch.gather(cdf, 1, inds_g).view(-1, n_sample, 2)
    bins_g = torch.gather(bins, 1, inds_g).view(-1, n_sample, 2)
    denom = cdf_g[..., 1] - cdf_g[..., 0]
    denom = torch.where(denom < eps, torch.ones_like(denom), denom)
    t = (u - cdf_g[..., 0]) / denom
    samples = bins_g[..., 0] + t * (bins_g[..., 1] - bins_g[..., 0])
    return torch.sort(samples, -1)[0]


def sample_pdf(bins, weights, u, eps=1e-5):
    """ray_helper.py:410-431"""
    weights = weights + eps
    pdf = weights / torch.sum(weights, -1, keepdim=True)
    cdf = torch.cumsum(pdf, -1)
    cdf = torch.cat([torch.zeros_like(cdf[..., :1]), cdf], -1)
    return sample_cdf(bins, cdf, u, eps)


def huber_loss(rgb, target, delta=0.1, weight=3000.0):
    """ImgLoss(Huber) (arcnerf/loss/img_loss.py:48-98): weight * mean over all entries."""
    d = (rgb - target).abs()
    loss = torch.where(d < delta, 0.5 / delta * d ** 2, d - 0.5 * delta)
    return weight * loss.mean()
