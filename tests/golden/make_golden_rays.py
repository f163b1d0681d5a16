"""Generates tests/golden/rays_golden.npz by calling the REFERENCE's own get_rays (arcnerf/render/ray_helper.py:12-119) on CPU
in the build container, for a small camera in every mode the device kernel covers.

    python tests/golden/make_golden_rays.py
"""
import os
import sys
import types
import warnings

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, '/root/reference')
sys.path.insert(0, os.path.join(ROOT, 'tests'))
sys.path.insert(1, ROOT)
p3d = types.ModuleType('pytorch3d'); tr = types.ModuleType('pytorch3d.transforms')
rc = types.ModuleType('pytorch3d.transforms.rotation_conversions')
for name in ('axis_angle_to_matrix', 'matrix_to_axis_angle', 'matrix_to_rotation_6d', 'rotation_6d_to_matrix'):
    setattr(rc, name, lambda *a, **k: None)
sys.modules.update({'pytorch3d': p3d, 'pytorch3d.transforms': tr, 'pytorch3d.transforms.rotation_conversions': rc})
warnings.filterwarnings('ignore')
from arcnerf.render.ray_helper import get_rays  # noqa: E402
import ray_cases as rcs  # noqa: E402


def main():
    out = {}
    for name, (W, H, K, c2w, kw) in rcs.cases().items():
        o, d, idx, r = get_rays(W, H, torch.from_numpy(K), torch.from_numpy(c2w), **kw)
        out[name + '/rays_o'], out[name + '/rays_d'] = o.numpy(), d.numpy()
        if r is not None:
            out[name + '/rays_r'] = r.numpy()
        if idx is not None:
            out[name + '/index'] = np.asarray(idx, np.int64)
    np.savez_compressed(os.path.join(ROOT, 'tests', 'golden', 'rays_golden.npz'), **out)
    print('wrote rays_golden.npz', {k: v.shape for k, v in out.items()})


if __name__ == '__main__':
    main()
