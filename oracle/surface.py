"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): numpy restatement of arcnerf/geometry/ray.py:403-601 (sphere_tracing,
secant_root_finding), one rounded fp32 operation at a time, every ray carried through every iteration (what the reference's masks
amount to).  Pinned: tests/golden/surface_golden.npz holds the outputs of the reference's own functions on seeded rays and
analytic fields (tests/golden/make_golden_surface.py); tests/test_oracle_surface.py compares.  The reference's
`if ~ind_low.sum() > 0` (ray.py:571) never holds, so only the low end of the secant bracket moves - reproduced."""
import numpy as np

F32 = np.float32


def _bounds(v, n):
    v = np.asarray(v, F32)
    return np.broadcast_to(v.reshape(-1) if v.size == n else v.reshape(1), (n, )).astype(F32)


def _pts(o, d, z):
    return (o + (d * z[:, None]).astype(F32)).astype(F32)


def sphere_tracing(o, d, sdf_func, near=0.0, far=10.0, n_iter=100, threshold=0.001):
    n = o.shape[0]
    nr, fr = _bounds(near, n), _bounds(far, n)
    z, sdf = nr.copy(), np.zeros(n, F32)
    mask, obj = np.ones(n, bool), np.zeros(n, bool)
    thr = F32(threshold)
    for _ in range(n_iter):
        valid = ~obj & mask
        if not valid.any():
            break
        new = sdf_func(_pts(o, d, z)).astype(F32)
        sdf = np.where(valid, new, sdf)
        if (np.abs(sdf) < thr).all():
            break
        obj |= np.abs(sdf) < thr
        upd = ~obj & mask
        z = np.where(upd, (z + sdf).astype(F32), z)
        mask &= ~(z > fr)
        mask &= ~(z < nr)
    z = np.where(z <= nr, F32(0), z)
    return z[:, None], _pts(o, d, z), mask


def linspace01(n):
    i = np.arange(n)
    dt = F32(1.0) / F32(n - 1) if n > 1 else F32(0)
    return np.where(i < n // 2, np.float64(dt) * i, 1.0 - np.float64(dt) * (n - 1 - i)).astype(F32)


def secant_root_finding(o, d, geo_func, near=0.0, far=10.0, n_step=128, n_iter=20, threshold=0.001, level=0.0, grad_dir='ascent'):
    n = o.shape[0]
    nr, fr = _bounds(near, n), _bounds(far, n)
    t = linspace01(n_step)[None]
    step = ((nr[:, None] * (F32(1) - t)).astype(F32) + (fr[:, None] * t).astype(F32)).astype(F32)
    pts = (o[:, None, :] + (d[:, None, :] * step[:, :, None]).astype(F32)).astype(F32).reshape(-1, 3)
    diff = (geo_func(pts).astype(F32).reshape(n, n_step) - F32(level)).astype(F32)
    if grad_dir == 'descent':
        diff = (diff * F32(-1)).astype(F32)
    not_occ = diff[:, 0] > 0
    sign = np.concatenate([np.sign((diff[:, :-1] * diff[:, 1:]).astype(F32)), np.ones((n, 1), F32)], -1)
    cost = sign * np.arange(n_step, 0, -1).astype(F32)
    index = cost.argmin(-1)
    min_cost = cost[np.arange(n), index]
    mask = not_occ & (min_cost < 0) & (diff[np.arange(n), index] > 0)
    i2 = np.minimum(index + 1, n_step - 1)
    z_high, g_high = step[np.arange(n), index], diff[np.arange(n), index]
    z_low, g_low = step[np.arange(n), i2].copy(), diff[np.arange(n), i2].copy()

    def mid():
        with np.errstate(all='ignore'):
            return ((((-g_low) * (z_high - z_low).astype(F32)).astype(F32) / (g_high - g_low).astype(F32)).astype(F32) + z_low).astype(F32)
    z_mid = mid()
    z_init = z_mid.copy()
    thr = F32(threshold)
    if mask.any():
        for i in range(n_iter):
            if i > 0 and (np.abs(z_init[mask] - z_mid[mask]) < thr).all():
                break
            g = (geo_func(_pts(o, d, z_mid)).astype(F32) - F32(level)).astype(F32)
            if grad_dir == 'descent':
                g = (g * F32(-1)).astype(F32)
            low = mask & (g < 0)
            z_low = np.where(low, z_mid, z_low)
            g_low = np.where(low, g, g_low)
            z_mid = np.where(mask, mid(), z_mid)         # the high end never moves (ray.py:571)
    z = np.where(mask, z_mid, fr)
    z = np.where(~not_occ, F32(0), z)
    z = np.where(z <= nr, F32(0), z).astype(F32)
    return z[:, None], _pts(o, d, z), mask
