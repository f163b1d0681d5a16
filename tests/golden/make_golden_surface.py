"""Outputs of the reference's own sphere_tracing / secant_root_finding (arcnerf/geometry/ray.py:403-601) on the seeded rays and
analytic fields of tests/surface_cases.py -> tests/golden/surface_golden.npz.  Run in the build container (needs /root/reference):
python tests/golden/make_golden_surface.py"""
import os
import sys
import types
import warnings

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, '/root/reference')
sys.path.insert(0, os.path.join(ROOT, 'tests'))
p3d = types.ModuleType('pytorch3d'); tr = types.ModuleType('pytorch3d.transforms')
rc = types.ModuleType('pytorch3d.transforms.rotation_conversions')
for name in ('axis_angle_to_matrix', 'matrix_to_axis_angle', 'matrix_to_rotation_6d', 'rotation_6d_to_matrix'):
    setattr(rc, name, lambda *a, **k: None)
sys.modules.update({'pytorch3d': p3d, 'pytorch3d.transforms': tr, 'pytorch3d.transforms.rotation_conversions': rc})
warnings.filterwarnings('ignore')
from arcnerf.geometry.ray import surface_ray_intersection  # noqa: E402
import surface_cases as sc  # noqa: E402

out = {}
o, d, near, far = sc.rays()
for name, (method, fname, kw) in sc.CASES.items():
    kw = dict(kw)
    for k in ('near', 'far'):
        if kw[k] == 'per_ray':
            kw[k] = torch.from_numpy(near if k == 'near' else far)
    f = sc.field(fname, torch)
    z, pts, mask = surface_ray_intersection(torch.from_numpy(o), torch.from_numpy(d), f, method=method, **kw)
    out[name + '/zvals'], out[name + '/pts'], out[name + '/mask'] = z.numpy(), pts.numpy(), mask.numpy()
    print(name, 'hit', int(mask.sum()), 'of', len(mask), 'z range', float(z.min()), float(z.max()))
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'surface_golden.npz'), **out)
