"""Seeded inputs for the ray-integration tests (shared by the golden generator, the CPU and the GPU tests)."""
import numpy as np

DENSE_CASES = ['plain', 'bkg_per_ray', 'bkg_single', 'white', 'inf_z', 'padded_tail', 'p33', 'p2']


def unit_dirs(n, seed):
    rng = np.random.RandomState(seed)
    d = rng.normal(size=(n, 3))
    return (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)


def dense_case(name):
    seed = DENSE_CASES.index(name) + 100
    rng = np.random.RandomState(seed)
    r, p = 37, 65
    if name == 'p33':
        r, p = 5, 33     # one interval past a 32-wide chunk boundary
    if name == 'p2':
        r, p = 9, 2      # the minimum row length reduce_empty_mask allows
    z = np.sort(rng.uniform(2.0, 6.0, size=(r, p)).astype(np.float32), axis=1)
    sigma = (rng.uniform(-2.0, 12.0, size=(r, p)) * (rng.uniform(size=(r, p)) < 0.7)).astype(np.float32)
    rad = rng.uniform(0, 1, size=(r, p, 3)).astype(np.float32)
    c = dict(sigma=sigma, rad=rad, zvals=z, add_inf_z=False, white_bkg=False, bkg=None)
    if name == 'bkg_per_ray':
        c['bkg'] = rng.uniform(0, 1, size=(r, 3)).astype(np.float32)
    if name == 'bkg_single':
        c['bkg'] = rng.uniform(0, 1, size=(1, 3)).astype(np.float32)
    if name == 'white':
        c['white_bkg'] = True
    if name == 'inf_z':
        c['add_inf_z'] = True
    if name == 'padded_tail':   # FgModel's padding: the tail repeats the last sample
        c['bkg'] = rng.uniform(0, 1, size=(r, 3)).astype(np.float32)
        for i in range(r):
            k = rng.randint(1, p + 1)
            z[i, k:] = z[i, k - 1]
            sigma[i, k:] = sigma[i, k - 1]
            rad[i, k:] = rad[i, k - 1]
    c['g_rgb'] = rng.normal(size=(r, 3)).astype(np.float32)
    c['g_depth'] = rng.normal(size=(r,)).astype(np.float32)
    c['g_mask'] = rng.normal(size=(r,)).astype(np.float32)
    return c


def alpha_case():
    rng = np.random.RandomState(7)
    r, p = 21, 40
    alpha = (rng.uniform(0, 0.6, size=(r, p)) * (rng.uniform(size=(r, p)) < 0.6)).astype(np.float32)
    return dict(alpha=alpha, sdf=rng.normal(size=(r, p)).astype(np.float32),
                rad=rng.uniform(0, 1, size=(r, p, 3)).astype(np.float32),
                zvals=np.sort(rng.uniform(1, 5, size=(r, p)).astype(np.float32), axis=1),
                bkg=rng.uniform(0, 1, size=(r, 3)).astype(np.float32),
                g_rgb=rng.normal(size=(r, 3)).astype(np.float32), g_w=rng.normal(size=(r, p)).astype(np.float32))


def sdf_case():
    rng = np.random.RandomState(8)
    r, p = 19, 31
    z = np.sort(rng.uniform(1, 5, size=(r, p + 1)).astype(np.float32), axis=1)
    return dict(sdf=rng.normal(scale=0.3, size=(r, p)).astype(np.float32), zvals=z,
                slope=-rng.uniform(0, 1, size=(r, p)).astype(np.float32), s=np.float32(20.0),
                g=rng.normal(size=(r, p)).astype(np.float32))


def pdf_case():
    rng = np.random.RandomState(9)
    r, nb = 23, 63
    bins = np.sort(rng.uniform(2, 6, size=(r, nb)).astype(np.float32), axis=1)
    w = (rng.uniform(0, 1, size=(r, nb - 1)) ** 4).astype(np.float32)
    return dict(bins=bins, weights=w, n_sample=128)
