"""Generates tests/golden/models_golden.npz by running the REFERENCE's own models on CPU in the build container:
FullModel(NeRF foreground + NeRFPP background) built by arcnerf.models.build_model from configs/models/nerfpp.yaml, and the
foreground-only NeRF of configs/models/nerf.yaml, with the seeded weights of tests/nerf_full_cases.py, jitter and density
noise switched off (rays.perturb False, rays.noise_std 0) so that the outputs are deterministic.

    python tests/golden/make_golden_models.py
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
p3d = types.ModuleType('pytorch3d'); tr = types.ModuleType('pytorch3d.transforms')
rc = types.ModuleType('pytorch3d.transforms.rotation_conversions')
for name in ('axis_angle_to_matrix', 'matrix_to_axis_angle', 'matrix_to_rotation_6d', 'rotation_6d_to_matrix'):
    setattr(rc, name, lambda *a, **k: None)
sys.modules.update({'pytorch3d': p3d, 'pytorch3d.transforms': tr, 'pytorch3d.transforms.rotation_conversions': rc})
warnings.filterwarnings('ignore')
from arcnerf.models import build_model  # noqa: E402
from common.utils.cfgs_utils import load_configs  # noqa: E402
import model_cases as mc  # noqa: E402
import nerf_full_cases as nc  # noqa: E402


def load(net_geo, net_rad, name, seed):
    gs, rs = nc.weights(name, seed)
    net_geo.load_state_dict({k: torch.from_numpy(v) for k, v in gs.items()})
    net_rad.load_state_dict({k: torch.from_numpy(v) for k, v in rs.items()})


def main():
    out = {}
    o, d = mc.rays()
    feed = {'rays_o': torch.from_numpy(o)[None], 'rays_d': torch.from_numpy(d)[None], 'rays_r': torch.zeros(1, o.shape[0], 1)}
    for tag, yaml in (('nerfpp', 'nerfpp.yaml'), ('nerf', 'nerf.yaml')):
        cfgs = load_configs('/root/reference/configs/models/' + yaml, None)
        cfgs.model.rays.perturb, cfgs.model.rays.noise_std = False, 0.0
        if tag == 'nerfpp':
            cfgs.model.background.rays.perturb, cfgs.model.background.rays.noise_std = False, 0.0
        model = build_model(cfgs, None)
        fg = model.fg_model
        load(fg.coarse_geo_net, fg.coarse_radiance_net, 'nerf', mc.SEEDS['fg_coarse'])
        load(fg.fine_geo_net, fg.fine_radiance_net, 'nerf', mc.SEEDS['fg_fine'])
        if tag == 'nerfpp':
            bk = model.bkg_model
            load(bk.coarse_geo_net, bk.coarse_radiance_net, 'nerfpp_bkg', mc.SEEDS['bkg_coarse'])
            load(bk.fine_geo_net, bk.fine_radiance_net, 'nerfpp_bkg', mc.SEEDS['bkg_fine'])
        with torch.no_grad():
            inf = model(dict(feed), inference_only=True)
            trn = model(dict(feed), inference_only=False)
        for k, v in inf.items():
            out['%s/inference/%s' % (tag, k)] = v[0].numpy()
        for k, v in trn.items():
            out['%s/train/%s' % (tag, k)] = v[0].numpy()
        out[tag + '/ray_cfgs'] = np.array(str({k: fg.get_ray_cfgs(k) for k in ('near', 'far', 'bounding_radius', 'n_sample', 'n_importance',
                                                                           'add_inf_z', 'white_bkg')}))
    np.savez_compressed(os.path.join(ROOT, 'tests', 'golden', 'models_golden.npz'), **out)
    print('wrote models_golden.npz:', {k: v.shape for k, v in out.items()})


if __name__ == '__main__':
    main()
