"""Generates tests/golden/nerf_full_golden.npz by IMPORTING the reference's own GeoNet / RadianceNet / FreqEmbedder from
/root/reference (build container only) at the FULL width of configs/models/nerf.yaml and nerfpp.yaml, loading the seeded
weights of tests/nerf_full_cases.py, and recording sigma, rgb and the gradients of  sum(w_sigma * sigma) + sum(w_rgb * rgb)
with respect to every parameter (biases in full, the first 8 rows of every weight matrix and its total, to keep the file small).

    python tests/golden/make_golden_nerf_full.py
"""
import os
import sys
import types

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
import warnings  # noqa: E402
warnings.filterwarnings('ignore')
from arcnerf.models.base_modules.geo_rad_model.linear_network_module import GeoNet, RadianceNet  # noqa: E402
from common.utils.cfgs_utils import dict_to_obj  # noqa: E402
import nerf_full_cases as nc  # noqa: E402


def main():
    out = {}
    for name, (d_x, nfx, nfv, W, D, skips, W_rad) in nc.CASES.items():
        geo = GeoNet(W=W, D=D, skips=skips, encoder=dict_to_obj(dict(type='FreqEmbedder', input_dim=d_x, n_freqs=nfx)), W_feat=W,
                     geometric_init=False)
        rad = RadianceNet(mode='vf', W=W_rad, D=1, encoder=dict_to_obj(dict(view=dict(type='FreqEmbedder', input_dim=3, n_freqs=nfv))),
                          W_feat_in=W)
        gs, rs_ = nc.weights(name)
        geo.load_state_dict({k: torch.from_numpy(v) for k, v in gs.items()})
        rad.load_state_dict({k: torch.from_numpy(v) for k, v in rs_.items()})
        x, dirs, w_s, w_c = (torch.from_numpy(a) for a in nc.inputs(name))
        sigma, feat = geo(x)
        rgb = rad(None, dirs, None, feat)
        ((sigma * w_s).sum() + (rgb * w_c).sum()).backward()
        out[name + '/sigma'], out[name + '/rgb'] = sigma.detach().numpy(), rgb.detach().numpy()
        for tag, net in (('geo', geo), ('rad', rad)):
            for k, p in net.named_parameters():
                g = p.grad.numpy()
                if k.endswith('bias'):
                    out['%s/grad/%s/%s' % (name, tag, k)] = g
                else:
                    out['%s/grad/%s/%s.rows8' % (name, tag, k)] = g[:8].copy()
                    out['%s/grad/%s/%s.colsum' % (name, tag, k)] = g.sum(0)
                    out['%s/grad/%s/%s.absmax' % (name, tag, k)] = np.array(np.abs(g).max())
    np.savez_compressed(os.path.join(ROOT, 'tests', 'golden', 'nerf_full_golden.npz'), **out)
    print('wrote nerf_full_golden.npz with', len(out), 'arrays')


if __name__ == '__main__':
    main()
