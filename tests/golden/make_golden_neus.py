"""Runs the REFERENCE's own Neus model (arcnerf/models/neus_model.py) on CPU with the configuration of tests/neus_cases.py and stores
its weights, the up-sampled depth samples (upsample_zvals, :106-171) and its rendered outputs -> tests/golden/neus_golden.npz.
    python tests/golden/make_golden_neus.py"""
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
from arcnerf.models import build_model  # noqa: E402
from common.utils.cfgs_utils import load_configs  # noqa: E402
import neus_cases as nc  # noqa: E402


def set_cfg(node, d):
    for k, v in d.items():
        if isinstance(v, dict):
            if not hasattr(node, k) or getattr(node, k) is None:
                setattr(node, k, types.SimpleNamespace())
            set_cfg(getattr(node, k), v)
        else:
            setattr(node, k, v)


def main():
    torch.manual_seed(7)
    cfgs = load_configs('/root/reference/configs/models/neus.yaml', None)
    set_cfg(cfgs.model, nc.CFG)
    model = build_model(cfgs, None)
    fg = model.fg_model
    out = {}
    for k, v in fg.state_dict().items():
        out['sd/' + k] = v.detach().numpy()
    o, d = nc.rays()
    ro, rd = torch.from_numpy(o), torch.from_numpy(d)
    with torch.no_grad():
        near, far, mask_rays = fg.get_near_far_from_rays({'rays_o': ro, 'rays_d': rd})
        zvals, mask_pts = fg.get_zvals_from_near_far(near, far, fg.get_ray_cfgs('n_sample'), True, rays_o=ro, rays_d=rd)
        z_up, m_up = fg.upsample_zvals(ro, rd, zvals, mask_pts, inference_only=True)
        res = model({'rays_o': ro[None], 'rays_d': rd[None], 'rays_r': torch.zeros(1, o.shape[0], 1)}, inference_only=True)
    out['near'], out['far'], out['mask_rays'] = near.numpy(), far.numpy(), (mask_rays.numpy() if mask_rays is not None else np.ones(len(o), bool))
    out['zvals'], out['z_up'] = zvals.numpy(), z_up.numpy()
    for k, v in res.items():
        if torch.is_tensor(v):
            out['out/' + k] = v[0].numpy()
    np.savez_compressed(os.path.join(ROOT, 'tests', 'golden', 'neus_golden.npz'), **out)
    print({k: v.shape for k, v in out.items() if not k.startswith('sd/')}, 'hit rays', int(out['mask_rays'].sum()))


if __name__ == '__main__':
    main()
