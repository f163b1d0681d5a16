"""Voxel-centre lattices of the reference's Volume (arcnerf/geometry/volume.py:225-232) -> tests/golden/volume_golden.npz.
Run in the build container (needs /root/reference): python tests/golden/make_golden_volume.py"""
import os
import sys

import types
import warnings

import numpy as np

sys.path.insert(0, '/root/reference')
# pytorch3d is not in this image; volume.py only imports the names
p3d = types.ModuleType('pytorch3d'); tr = types.ModuleType('pytorch3d.transforms')
rc = types.ModuleType('pytorch3d.transforms.rotation_conversions')
for name in ('axis_angle_to_matrix', 'matrix_to_axis_angle', 'matrix_to_rotation_6d', 'rotation_6d_to_matrix'):
    setattr(rc, name, lambda *a, **k: None)
sys.modules.update({'pytorch3d': p3d, 'pytorch3d.transforms': tr, 'pytorch3d.transforms.rotation_conversions': rc})
warnings.filterwarnings('ignore')
from arcnerf.geometry.volume import Volume  # noqa: E402

out = {}
for tag, kw in [('cube16', dict(n_grid=16, side=1.5)), ('box9', dict(n_grid=9, origin=(0.1, -0.2, 0.3), xyz_len=(2.0, 1.0, 3.0))),
                ('one', dict(n_grid=1, side=2.0)),
                ('box67', dict(n_grid=67, origin=(-0.3, 0.2, 0.05), xyz_len=(1.7, 2.9, 0.9)))]:
    vol = Volume(**kw)
    rng = vol.get_range().numpy()
    out['%s_lo' % tag], out['%s_hi' % tag] = rng[:, 0].astype(np.float32), rng[:, 1].astype(np.float32)
    out['%s_n' % tag] = np.int32(kw['n_grid'])
    out['%s_pts' % tag] = vol.get_volume_pts().numpy().astype(np.float32)
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'volume_golden.npz'), **out)
print({k: v.shape for k, v in out.items()})
