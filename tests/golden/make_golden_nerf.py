"""Generates tests/golden/nerf_golden.npz by IMPORTING the reference's own modules from /root/reference (build container
only): FreqEmbedder (freq_encoder.py), GeoNet and RadianceNet (linear_network_module.py) at reduced width so that the
fixture stays small.  The state dicts are stored, so the mirrors are loaded with the reference's weights.

    python tests/golden/make_golden_nerf.py
"""
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, '/root/reference')
p3d = types.ModuleType('pytorch3d'); tr = types.ModuleType('pytorch3d.transforms')
rc = types.ModuleType('pytorch3d.transforms.rotation_conversions')
for name in ('axis_angle_to_matrix', 'matrix_to_axis_angle', 'matrix_to_rotation_6d', 'rotation_6d_to_matrix'):
    setattr(rc, name, lambda *a, **k: None)
sys.modules.update({'pytorch3d': p3d, 'pytorch3d.transforms': tr, 'pytorch3d.transforms.rotation_conversions': rc})
import warnings  # noqa: E402
warnings.filterwarnings('ignore')
from arcnerf.models.base_modules.encoding.freq_encoder import FreqEmbedder  # noqa: E402
from arcnerf.models.base_modules.geo_rad_model.linear_network_module import GeoNet, RadianceNet  # noqa: E402
from common.utils.cfgs_utils import dict_to_obj  # noqa: E402


def main():
    torch.manual_seed(0)
    out = {}
    x = torch.rand((257, 3)) * 2 - 1
    out['x'] = x.numpy()
    for tag, kw in (('log10', dict(n_freqs=10)), ('lin4', dict(n_freqs=4, log_sampling=False)),
                    ('log6_noinput', dict(n_freqs=6, include_input=False)), ('none', dict(n_freqs=0))):
        enc = FreqEmbedder(3, **kw)
        xr = x.clone().requires_grad_(True)
        y = enc(xr)
        w = torch.rand(y.shape, generator=torch.Generator().manual_seed(len(tag)))
        (y * w).sum().backward()
        out['freq_%s' % tag], out['freq_%s_w' % tag], out['freq_%s_gx' % tag] = y.detach().numpy(), w.numpy(), xr.grad.numpy()
    # vanilla-NeRF shaped networks at reduced width: D=4, W=64, skip at 2, 63-dim position / 27-dim view encodings
    geo = GeoNet(W=64, D=4, skips=[2], encoder=dict_to_obj(dict(type='FreqEmbedder', input_dim=3, n_freqs=10)), W_feat=64,
                 geometric_init=False)
    rad = RadianceNet(mode='vf', W=32, D=1, encoder=dict_to_obj(dict(view=dict(type='FreqEmbedder', input_dim=3, n_freqs=4))),
                      W_feat_in=64)
    for name, net in (('geo', geo), ('rad', rad)):
        for k, v in net.state_dict().items():
            out['%s_sd/%s' % (name, k)] = v.numpy()
    dirs = torch.randn((257, 3))
    sigma, feat = geo(x)
    rgb = rad(None, dirs, None, feat)
    out.update(dirs=dirs.numpy(), sigma=sigma.detach().numpy(), feat=feat.detach().numpy(), rgb=rgb.detach().numpy())
    # geometric init variant with softplus and forward_with_grad (the SDF-model use of the same class)
    geo2 = GeoNet(W=64, D=3, skips=[1], encoder=dict_to_obj(dict(type='FreqEmbedder', input_dim=3, n_freqs=6)), W_feat=8,
                  geometric_init=True, radius_init=0.5, norm_skip=True, skip_reduce_output=True,
                  act_cfg=dict_to_obj(dict(type='Softplus', beta=100)))
    for k, v in geo2.state_dict().items():
        out['geo2_sd/%s' % k] = v.numpy()
    sdf, f2, grad = geo2.forward_with_grad(x.clone())
    out.update(sdf=sdf.detach().numpy(), feat2=f2.detach().numpy(), sdf_grad=grad.detach().numpy())
    # the SDF network of the NeuS-NGP configuration (nerf_lego_neus_ngp.yaml geometry: W 64, D 1, no bias, weight norm, softplus 100,
    # geometric init) on a 32-dim input standing in for the hash-grid features, with the second-order pass of the eikonal loss
    torch.manual_seed(3)
    geo3 = GeoNet(W=64, D=1, skips=[], encoder=dict_to_obj(dict(type='FreqEmbedder', input_dim=32, n_freqs=0)), W_feat=16, use_bias=False,
                  geometric_init=True, radius_init=0.5, weight_norm=True, skip_reduce_output=True, norm_skip=True,
                  act_cfg=dict_to_obj(dict(type='Softplus', beta=100)))
    for k, v in geo3.state_dict().items():
        out['geo3_sd/%s' % k] = v.numpy()
    x3 = torch.randn((300, 32)) * 0.3
    w1, w2, w3 = torch.randn((300, 1)), torch.randn((300, 16)), torch.randn((300, 32))
    sdf3, feat3, grad3 = geo3.forward_with_grad(x3.clone())
    loss = ((grad3.norm(dim=-1) - 1.0) ** 2).mean() + (sdf3 * w1).sum() + (feat3 * w2).sum() + (grad3 * w3).sum()
    loss.backward()
    out.update(x3=x3.numpy(), w1_3=w1.numpy(), w2_3=w2.numpy(), w3_3=w3.numpy(), sdf3=sdf3.detach().numpy(), feat3=feat3.detach().numpy(),
               grad3=grad3.detach().numpy(), loss3=np.array(float(loss)))
    for k, p in geo3.named_parameters():
        out['geo3_grad/%s' % k] = p.grad.numpy()
    np.savez_compressed(os.path.join(ROOT, 'tests', 'golden', 'nerf_golden.npz'), **out)
    print('wrote nerf_golden.npz with', len(out), 'arrays')


if __name__ == '__main__':
    main()
