"""Generates tests/golden/zvals_golden.npz by IMPORTING the reference's own functions from /root/reference (build
container only): arcnerf.geometry.ray.sphere_ray_intersection, arcnerf.render.ray_helper.{get_near_far_from_rays,
get_zvals_from_near_far, get_zvals_from_near_far_fix_step, get_zvals_outside_sphere, perturb_interval} and
arcnerf.geometry.volume.Volume (opacity-field update).  The reference draws its jitter with torch.rand; the generator
swaps torch.rand for a function returning the stored uniforms while the reference code runs, so that the golden outputs
are functions of the stored inputs only.

    python tests/golden/make_golden_zvals.py
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
from arcnerf.geometry.ray import sphere_ray_intersection  # noqa: E402
from arcnerf.geometry.volume import Volume  # noqa: E402
from arcnerf.render import ray_helper as rh  # noqa: E402


class FixedRand:
    """context manager: torch.rand(...) returns the next stored tensor"""

    def __init__(self, *tensors):
        self.queue = list(tensors)

    def __enter__(self):
        self.real = torch.rand
        torch.rand = lambda *a, **k: self.queue.pop(0).clone()

    def __exit__(self, *exc):
        torch.rand = self.real


def main():
    g = torch.Generator().manual_seed(11)
    n = 64
    # rays: origins on a shell of radius 3 looking roughly at the origin, plus a few inside / tangent / missing ones
    o = torch.randn((n, 3), generator=g)
    o = o / o.norm(dim=1, keepdim=True) * 3.0
    d = -o + 0.6 * torch.randn((n, 3), generator=g)
    d = d / d.norm(dim=1, keepdim=True)
    o[:8] = 0.3 * torch.randn((8, 3), generator=g)                  # inside the unit sphere
    o[8], d[8] = torch.tensor([1.0, -3.0, 0.0]), torch.tensor([0.0, 1.0, 0.0])   # tangent to r = 1
    o[9], d[9] = torch.tensor([2.5, -3.0, 0.0]), torch.tensor([0.0, 1.0, 0.0])   # misses r <= 2
    out = dict(rays_o=o.numpy(), rays_d=d.numpy())
    radius = torch.tensor([1.0, 1.5, 2.0, 4.0])
    near, far, pts, mask = sphere_ray_intersection(o.clone(), d.clone(), radius)
    out.update(sph_radius=radius.numpy(), sph_near=near.numpy(), sph_far=far.numpy(), sph_pts=pts.numpy(), sph_mask=mask.numpy())

    bounds = torch.stack([1.0 + torch.rand(n, generator=g), 4.0 + 2 * torch.rand(n, generator=g)], 1)
    out['bounds'] = bounds.numpy()
    for tag, kw in (('nf_radius', dict(bounding_radius=2.0)), ('nf_bounds', dict(bounds=bounds.clone())),
                    ('nf_bounds_radius', dict(bounds=bounds.clone(), bounding_radius=2.0)),
                    ('nf_hard', dict(near_hardcode=2.0, far_hardcode=6.0)),
                    ('nf_near_hard', dict(bounding_radius=3.5, near_hardcode=0.5))):
        a, b = rh.get_near_far_from_rays(o.clone(), d.clone(), **kw)
        out[tag + '_near'], out[tag + '_far'] = a.numpy(), b.numpy()

    near, far = bounds[:, 0:1].clone(), bounds[:, 1:2].clone()
    for n_pts in (64, 7):
        u = torch.rand((n, n_pts), generator=g)
        out['u_%d' % n_pts] = u.numpy()
        for inc in (True, False):
            for inv in (False, True):
                key = 'zv_%d_%d_%d' % (n_pts, inc, inv)
                out[key] = rh.get_zvals_from_near_far(near, far, n_pts, inclusive=inc, inverse_linear=inv).numpy()
                with FixedRand(u):
                    out[key + '_p'] = rh.get_zvals_from_near_far(near, far, n_pts, inclusive=inc, inverse_linear=inv,
                                                                 perturb=True).numpy()
    # fixed step: some rays run out of range before n_pts samples, some do not
    near_f = 0.5 + torch.rand((n, 1), generator=g)
    far_f = near_f + 0.2 + 2.5 * torch.rand((n, 1), generator=g)
    out['fix_near'], out['fix_far'] = near_f.numpy(), far_f.numpy()
    for n_pts, fix_t in ((48, 0.05), (33, 0.011)):
        u = torch.rand((n, n_pts), generator=g)
        out['fix_u_%d' % n_pts] = u.numpy()
        for inc in (True, False):
            with FixedRand(u):
                z, m = rh.get_zvals_from_near_far_fix_step(near_f.clone(), far_f.clone(), fix_t, n_pts, inclusive=inc)
            out['fix_z_%d_%d' % (n_pts, inc)], out['fix_m_%d_%d' % (n_pts, inc)] = z.numpy(), m.numpy()
    # NeRF++ shells
    n_pts = 32
    z, sr = rh.get_zvals_outside_sphere(o.clone(), d.clone(), n_pts, 1.5)
    out['shell_z'], out['shell_r'] = z.numpy(), sr.numpy()
    u = torch.rand((1, n_pts), generator=g)
    out['shell_u'] = u.numpy()
    with FixedRand(u):
        z, sr = rh.get_zvals_outside_sphere(o.clone(), d.clone(), n_pts, 1.5, perturb=True)
    out['shell_z_p'], out['shell_r_p'] = z.numpy(), sr.numpy()
    # opacity field update of the volume bound
    vol = Volume(n_grid=8, side=2.0)
    vol.set_up_voxel_bitfield(init_occ=True)
    vol.set_up_voxel_opafield()
    vol.opafield.copy_(torch.rand((8, 8, 8), generator=g) * 0.02 - 0.002)       # a few negative (frozen) voxels
    idx = torch.unique(torch.randint(0, 8, (200, 3), generator=g), dim=0)
    opa = torch.rand((idx.shape[0], ), generator=g) * 0.03
    out.update(opa_field0=vol.opafield.clone().numpy(), opa_idx=idx.numpy(), opa_new=opa.numpy())
    vol.update_opafield_by_voxel_idx(idx, opa, ema=0.95)
    vol.update_bitfield_by_opafield(threshold=0.01, ops='overwrite')
    out.update(opa_field1=vol.opafield.numpy(), opa_bits1=vol.get_voxel_bitfield().numpy())
    np.savez_compressed(os.path.join(ROOT, 'tests', 'golden', 'zvals_golden.npz'), **out)
    print('wrote zvals_golden.npz with', len(out), 'arrays')


if __name__ == '__main__':
    main()
