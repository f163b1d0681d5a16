"""Seeded rays and analytic fields shared by the surface-intersection goldens, the CPU oracle test and the GPU test.  Every field
is written with elementwise + - * / sqrt only, evaluated the same way by numpy and torch (CPU or CUDA) up to the last bit of
the square root / division."""
import numpy as np

CASES = {
    # name: (method, field, kwargs)
    'trace_sphere': ('sphere_tracing', 'sphere', dict(near=0.0, far=6.0, n_iter=40, threshold=0.001)),
    'trace_two_spheres_near': ('sphere_tracing', 'two_spheres', dict(near=0.5, far=5.0, n_iter=60, threshold=0.0005)),
    'trace_per_ray_bounds': ('sphere_tracing', 'sphere', dict(near='per_ray', far='per_ray', n_iter=30, threshold=0.002)),
    'secant_sphere': ('secant_root_finding', 'sphere', dict(near=0.0, far=6.0, n_step=64, n_iter=20, threshold=0.001)),
    'secant_two_spheres': ('secant_root_finding', 'two_spheres', dict(near=0.2, far=5.0, n_step=128, n_iter=8, threshold=1e-6)),
    'secant_density': ('secant_root_finding', 'density', dict(near=0.0, far=6.0, n_step=96, n_iter=12, threshold=0.0005, level=50.0,
                                                             grad_dir='descent')),
    'secant_per_ray_bounds': ('secant_root_finding', 'sphere', dict(near='per_ray', far='per_ray', n_step=48, n_iter=10, threshold=0.001)),
}
N_RAYS = 257


def rays(seed=0, n=N_RAYS):
    rs = np.random.RandomState(seed)
    o = rs.normal(size=(n, 3))
    o = o / np.linalg.norm(o, axis=1, keepdims=True) * 3.0
    o[:5] *= 0.1                                            # a few origins inside the objects
    d = -o / np.linalg.norm(o, axis=1, keepdims=True) + 0.18 * rs.normal(size=(n, 3))
    d = d / np.linalg.norm(d, axis=1, keepdims=True)
    near = rs.uniform(0.0, 1.5, (n, 1))
    far = near + rs.uniform(1.0, 4.0, (n, 1))
    return o.astype(np.float32), d.astype(np.float32), near.astype(np.float32), far.astype(np.float32)


def field(name, xp):
    """xp: numpy or torch; returns f(pts (N,3)) -> (N,)"""
    def sq(p, c):
        x, y, z = p[:, 0] - c[0], p[:, 1] - c[1], p[:, 2] - c[2]
        return x * x + y * y + z * z
    if name == 'sphere':
        return lambda p: xp.sqrt(sq(p, (0.1, -0.05, 0.0))) - 0.6
    if name == 'two_spheres':
        return lambda p: xp.minimum(xp.sqrt(sq(p, (0.45, 0.0, 0.1))) - 0.4, xp.sqrt(sq(p, (-0.4, 0.2, -0.1))) - 0.5)
    if name == 'density':
        return lambda p: 120.0 / (1.0 + 9.0 * sq(p, (0.0, 0.0, 0.0)))
    raise KeyError(name)
