"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): CPU restatement of the depth-sample helpers around the march.

Follows arcnerf/geometry/ray.py:180-255 (sphere_ray_intersection), arcnerf/render/ray_helper.py:175-228
(get_near_far_from_rays), :231-263 (get_zvals_from_near_far), :266-315 (get_zvals_from_near_far_fix_step), :318-358
(get_zvals_outside_sphere), :361-407 (perturb_interval, perturb_interval_with_mask), arcnerf/models/nerf_model.py:93-117
(upsample_zvals) and the opacity-field rules of arcnerf/geometry/volume.py:983-1018.  Random draws are arguments
(`u`), so the functions are deterministic.  Pinned against the reference's own functions by
tests/golden/make_golden_zvals.py -> tests/golden/zvals_golden.npz (tests/test_oracle_zvals.py).
"""
import torch


def _zero_small(t, tol=1e-5):
    return torch.where(t.abs() < tol, torch.zeros_like(t), t)


def sphere_ray_intersection(rays_o, rays_d, radius, origin=(0.0, 0.0, 0.0)):
    """-> near, far (N, Nr), pts (N, Nr, 2, 3), mask (N, Nr)"""
    n, nr = rays_o.shape[0], radius.shape[0]
    o = rays_o[:, None, :].expand(n, nr, 3)
    d = rays_d[:, None, :].expand(n, nr, 3)
    r = radius[None, :].expand(n, nr)
    oc = torch.tensor(origin, dtype=rays_o.dtype)[None, None, :] - o
    z_half = _zero_small((oc * d).sum(-1))
    inside = torch.norm(oc, dim=-1) <= r
    mask = (z_half > 0) | inside
    d2 = _zero_small((oc * oc).sum(-1) - z_half * z_half)
    mask = mask & (d2 >= 0)
    z_off = _zero_small(r ** 2 - d2)
    mask = mask & (z_off >= 0)
    z_off = torch.sqrt(z_off.clamp_min(0.0))
    near = (z_half - z_off).clamp_min(0.0)
    far = (z_half + z_off).clamp_min(0.0)
    near = torch.where(mask, near, torch.zeros_like(near))
    far = torch.where(mask, far, torch.zeros_like(far))
    pts = torch.stack([o + d * near[..., None], o + d * far[..., None]], dim=2)
    return near, far, pts, mask


def get_near_far_from_rays(rays_o, rays_d, bounds=None, near_hardcode=None, far_hardcode=None, bounding_radius=None):
    n = rays_o.shape[0]
    if near_hardcode is None or far_hardcode is None:
        if bounds is None:
            near, far, _, _ = sphere_ray_intersection(rays_o, rays_d, torch.tensor([bounding_radius], dtype=rays_o.dtype))
        else:
            near, far = bounds[:, 0:1].clone(), bounds[:, 1:2].clone()
            if bounding_radius is not None:
                _, fb, _, _ = sphere_ray_intersection(rays_o, rays_d, torch.tensor([bounding_radius], dtype=rays_o.dtype))
                far = torch.where(far > fb, fb, far)
        if near_hardcode is not None:
            near = near * 0 + near_hardcode
        if far_hardcode is not None:
            far = far * 0 + far_hardcode
    else:
        near = torch.ones((n, 1), dtype=rays_o.dtype) * near_hardcode
        far = torch.ones((n, 1), dtype=rays_o.dtype) * far_hardcode
    far = torch.where(far <= near, near + 1e-5, far)
    return near, far


def perturb_interval(vals, u):
    mids = .5 * (vals[..., 1:] + vals[..., :-1])
    upper = torch.cat([mids, vals[..., -1:]], -1)
    lower = torch.cat([vals[..., :1], mids], -1)
    return lower + (upper - lower) * u


def get_zvals_from_near_far(near, far, n_pts, inclusive=True, inverse_linear=False, u=None):
    if inclusive:
        t = torch.linspace(0.0, 1.0, n_pts, dtype=near.dtype)
    else:
        t = torch.linspace(0.0, 1.0, n_pts + 2, dtype=near.dtype)[1:-1]
    if inverse_linear:
        z = 1.0 / (1.0 / (near + 1e-8) * (1.0 - t) + 1.0 / (far + 1e-8) * t)
    else:
        z = near + (far - near) * t
    return z if u is None else perturb_interval(z, u)


def get_zvals_from_near_far_fix_step(near, far, fix_t, n_pts, inclusive=True, u=None):
    n = near.shape[0]
    z = torch.ones((n, n_pts), dtype=near.dtype) * (near if inclusive else near + fix_t)
    z = z + torch.arange(0, n_pts, 1)[None] * fix_t
    z = torch.min(torch.max(z, near), far)
    same = torch.cat([torch.zeros((n, 1), dtype=torch.bool), (z[:, 1:] - z[:, :-1]) == 0.0], dim=1)
    mask = ~same
    if u is None:
        return z, mask
    pert = perturb_interval(z, u)
    z = torch.where(mask, pert, z)
    last = mask.sum(dim=1) - 1
    last_val = z[torch.arange(n), last][:, None]
    return torch.min(torch.max(z, z[:, 0:1]), last_val), mask


def get_zvals_outside_sphere(rays_o, rays_d, n_pts, radius, u=None):
    t = torch.linspace(0.0, 1.0, n_pts + 2, dtype=rays_o.dtype)[1:-1]
    sr = radius / torch.flip(t, dims=[-1])
    if u is not None:
        sr = perturb_interval(sr[None], u)[0]
    _, z, _, _ = sphere_ray_intersection(rays_o, rays_d, sr)
    return z, sr


def update_opafield(opafield, bitfield, voxel_idx, opacity, ema, threshold):
    """Volume.update_opafield_by_voxel_idx + update_bitfield_by_opafield(ops='overwrite'): -> new opafield, bitfield"""
    opa = opafield.clone()
    cur = opa[voxel_idx[:, 0], voxel_idx[:, 1], voxel_idx[:, 2]]
    new = torch.max(cur * ema, opacity) if ema is not None else opacity
    opa[voxel_idx[:, 0], voxel_idx[:, 1], voxel_idx[:, 2]] = torch.where(cur >= 0, new, cur)
    thres = min(float(opa.clamp(min=0).mean()), threshold)
    return opa, opa >= thres
