"""Seeded input cases shared by the golden generator, the CPU tests and the GPU parity tests."""
import numpy as np

from arcnerf_b200 import synth

# The reference's own 8-ray AABB table (tests/tests_arcnerf/tests_geometry/tests_ray.py:329-349):
# inside / outside / tangent / on-surface origins against the side-2 cube.
AABB_TABLE_O = np.array([[0.5, 0.3, 0], [-0.5, -0.1, -0.2], [0.5, 2.0, 1.0], [1.8, 1.4, 2.1], [-0.5, -0.5, 2.0],
                         [0.2, 0, 1], [0, 0.2, 1], [0, -0.2, 1]], np.float32)
AABB_TABLE_D = np.array([[-1.0, -1.0, 0], [-1.0, -1.0, -1.0], [0, -1.0, 0.0], [-0.9, -0.7, -1.0], [-1, -1, 1.0],
                         [0, 0, 1], [0, 0, -1], [1, 0.5, 0]], np.float32)


def aabb_table():
    d = AABB_TABLE_D / np.linalg.norm(AABB_TABLE_D, axis=1, keepdims=True)
    return AABB_TABLE_O.copy(), d.astype(np.float32)


def rays_towards_origin(n, seed, radius=3.0, jitter=0.6):
    """Random origins on a sphere, directions towards a jittered point near the origin (the reference's model
    tests use rays pointing at the origin: tests/tests_arcnerf/tests_models/__init__.py:97-119)."""
    rng = np.random.RandomState(seed)
    o = rng.normal(size=(n, 3))
    o = o / np.linalg.norm(o, axis=1, keepdims=True) * radius
    tgt = rng.uniform(-jitter, jitter, size=(n, 3))
    d = tgt - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return o.astype(np.float32), d.astype(np.float32)


def march_case(name):
    """Returns dict(o, d, aabb(2,3), vol(n,n,n) bool, n_grid, n_pts, dt, near_distance)."""
    if name == 'small32':
        o, d = rays_towards_origin(256, 1)
        n = 32
        vol = synth.occupancy_volume('lego', n=n, seed=3)
        return dict(o=o, d=d, aabb=synth.aabb_of_side(2.0), vol=vol, n_grid=n, n_pts=128,
                    dt=np.float32(2 * np.sqrt(3.0) / 128), near_distance=0.2)
    if name == 'core50_64':
        o, d = rays_towards_origin(192, 2, jitter=0.9)
        n = 64
        vol = synth.occupancy_volume('core50', n=n, seed=5)
        return dict(o=o, d=d, aabb=synth.aabb_of_side(2.0), vol=vol, n_grid=n, n_pts=96,
                    dt=np.float32(2 * np.sqrt(3.0) / 256), near_distance=0.0)
    if name == 'offcentre':  # non-cubic, off-centre box; exercises the n_grid*pos quirk of the skip distance
        o, d = rays_towards_origin(128, 7, radius=4.0, jitter=0.5)
        o += np.array([0.3, -0.2, 0.1], np.float32)
        n = 32
        vol = synth.occupancy_volume('sphere', n=n)
        aabb = np.array([[-0.9, -1.3, -0.8], [1.5, 0.9, 1.0]], np.float32)
        return dict(o=o, d=d, aabb=aabb, vol=vol, n_grid=n, n_pts=64, dt=np.float32(0.02), near_distance=0.1)
    if name == 'full_cap':  # fully occupied: every ray hits the n_pts cap
        o, d = rays_towards_origin(64, 9)
        n = 16
        return dict(o=o, d=d, aabb=synth.aabb_of_side(2.0), vol=np.ones((n, n, n), bool), n_grid=n, n_pts=48,
                    dt=np.float32(0.01), near_distance=0.0)
    if name == 'empty':
        o, d = rays_towards_origin(64, 10)
        n = 16
        return dict(o=o, d=d, aabb=synth.aabb_of_side(2.0), vol=np.zeros((n, n, n), bool), n_grid=n, n_pts=32,
                    dt=np.float32(0.02), near_distance=0.0)
    if name == 'ngp4096':  # BASELINE configs[1] shape: 4096 Blender-shaped rays, 128^3, 1024 samples
        o, d = synth.train_ray_batch(4096, seed=0)
        return dict(o=o, d=d, aabb=synth.aabb_of_side(2.0), vol=synth.occupancy_volume('lego'), n_grid=128,
                    n_pts=1024, dt=np.float32(2 * np.sqrt(3.0) / 1024), near_distance=0.2)
    raise KeyError(name)


MARCH_GOLDEN_CASES = ['small32', 'core50_64', 'offcentre', 'full_cap', 'empty']


def multivol_case(inclusive):
    o, d = rays_towards_origin(128, 21, radius=0.3, jitter=0.2)   # cameras inside the inner volume
    rng = np.random.RandomState(22)
    d = rng.normal(size=d.shape).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    n, n_cascade = 16, 4
    n_level = n_cascade if inclusive else n_cascade - 1
    bits = (rng.uniform(size=(n ** 3 // 8 * n_level,)) < 0.35)
    bits = (rng.randint(0, 256, size=bits.shape) * bits).astype(np.uint8)
    min_aabb = synth.aabb_of_side(2.0)
    aabb = synth.aabb_of_side(2.0 * 2 ** (n_cascade - 1))
    return dict(o=o, d=d, min_aabb=min_aabb, aabb=aabb, bits=bits, n_grid=n, n_cascade=n_cascade, n_pts=64,
                cone=np.float32(1.0 / 256), smin=np.float32(0.02), smax=np.float32(0.5), near_distance=0.05,
                inclusive=inclusive)
