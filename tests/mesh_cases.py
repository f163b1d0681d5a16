"""Seeded volumes shared by the CPU and GPU mesh tests."""
import numpy as np


def blobs_volume(shape, seed=0):
    """smooth random field (sum of Gaussians) + a little noise, level chosen so the surface exists -> vol, voxel_size, volume_len, level"""
    rng = np.random.default_rng(seed)
    g = np.stack(np.meshgrid(*[np.linspace(-1, 1, s) for s in shape], indexing='ij'), -1)
    vol = np.zeros(shape)
    for _ in range(6):
        c, w = rng.uniform(-0.7, 0.7, 3), rng.uniform(0.15, 0.5)
        vol += np.exp(-((g - c) ** 2).sum(-1) / (w * w))
    vol += 0.05 * rng.standard_normal(shape)
    vol = (vol * 40).astype(np.float32)
    vl = (2.0, 1.5, 3.0)
    vs = tuple(vl[a] / shape[a] for a in range(3))
    return vol, vs, vl, float(np.median(vol) + 5.0)


def sphere_volume(n, r):
    ax = (np.arange(n) + 0.5) * (2.0 / n) - 1.0
    g = np.stack(np.meshgrid(ax, ax, ax, indexing='ij'), -1)
    vol = ((r - np.linalg.norm(g, axis=-1)) * 50 + 50).astype(np.float32)
    return vol, (2.0 / n, ) * 3, (2.0, ) * 3


def closed_and_oriented(v, f, manifold=True):
    """the boundary of the triangle chain vanishes: every directed edge is matched by its reverse, with multiplicity.  On smooth
    fields every directed edge also occurs once (manifold); on white noise two fan diagonals of neighbouring cells may coincide
    in the plane of an ambiguous face, which keeps the surface closed but uses the edge twice."""
    e = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]]).astype(np.int64)
    key, rev = e[:, 0] * len(v) + e[:, 1], e[:, 1] * len(v) + e[:, 0]
    if manifold:
        assert len(np.unique(key)) == len(key), 'a directed edge is used twice'
    assert np.array_equal(np.sort(key), np.sort(rev)), 'an edge has no opposite: the surface is open or inconsistently wound'
    assert f.min() >= 0 and f.max() < len(v) and len(np.unique(f)) == len(v)
