"""Generates tests/golden/render_golden.npz by IMPORTING the reference's own Python functions from
/root/reference (build container only; the GPU box has no /root/reference):

    arcnerf.render.ray_helper.{ray_marching, alpha_to_weights, sample_pdf}   (ray_helper.py:410-620)
    arcnerf.models.neus_model.sdf_to_alpha                                   (neus_model.py:242-265)
    arcnerf.models.base_modules.encoding.sh_encoder.SHEmbedder (torch path)  (sh_encoder.py:126-199)
    arcnerf.geometry.ray.aabb_ray_intersection (torch path)                  (ray.py:258-350)

The absent `pytorch3d` (only four rotation helpers, never used on this path) is stubbed, as SURVEY.md §7 notes.
Inputs follow the reference's own fixtures: the synthetic trapezoid-density ray of
tests/tests_arcnerf/tests_render/tests_ray_helper.py:68-86 plus seeded random rays.

    python tests/golden/make_golden_torch.py
"""
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
sys.path.insert(0, '/root/reference')

# stub for the absent pytorch3d import of arcnerf/geometry/transformation.py:6-8
p3d = types.ModuleType('pytorch3d'); tr = types.ModuleType('pytorch3d.transforms')
rc = types.ModuleType('pytorch3d.transforms.rotation_conversions')
for name in ('axis_angle_to_matrix', 'matrix_to_axis_angle', 'matrix_to_rotation_6d', 'rotation_6d_to_matrix'):
    setattr(rc, name, lambda *a, **k: None)
sys.modules.update({'pytorch3d': p3d, 'pytorch3d.transforms': tr, 'pytorch3d.transforms.rotation_conversions': rc})

import warnings  # noqa: E402
warnings.filterwarnings('ignore')
from arcnerf.render.ray_helper import ray_marching, alpha_to_weights, sample_pdf  # noqa: E402
from arcnerf.models.neus_model import sdf_to_alpha  # noqa: E402
from arcnerf.models.base_modules.encoding.sh_encoder import SHEmbedder  # noqa: E402

import render_cases  # noqa: E402


def trapezoid_ray(n_pts=128, near=2.0, far=6.0, append=False):
    """tests_ray_helper.py:68-86"""
    offset = int(n_pts / 3.0)
    level = 20
    sigma = [0.0 for _ in range(offset)]
    sigma.extend([(i + 1) * level for i in range(int(offset / 5.0) + 1)])
    sigma.extend([(int(offset / 5.0) + 1) * level for _ in range(3 * int((offset / 5.0) + 1))])
    sigma.extend([(int(offset / 5.0) + 1) * level - (i + 1) * level for i in range(int(offset / 5.0) + 1)])
    sigma.extend([0.0 for _ in range(offset)])
    sigma = np.array(sigma[:n_pts])[None, :]
    zvals = np.linspace(near, far, n_pts)[None, :]
    if append:
        zvals = np.concatenate([zvals, np.repeat(zvals[:, -1:], int(n_pts / 8), 1)], axis=-1)
        sigma = np.concatenate([sigma, np.repeat(sigma[:, -1:], int(n_pts / 8), 1)], axis=-1)
    return sigma.astype(np.float32), zvals.astype(np.float32)


def main():
    out = {}
    for append in (False, True):
        sigma, zvals = trapezoid_ray(append=append)
        rad = np.random.RandomState(0).uniform(0, 1, size=sigma.shape + (3,)).astype(np.float32)
        for add_inf_z in (False, True):
            for white in (False, True):
                r = ray_marching(torch.from_numpy(sigma), torch.from_numpy(rad), torch.from_numpy(zvals),
                                 add_inf_z=add_inf_z, white_bkg=white)
                key = 'trap_app%d_inf%d_white%d_' % (append, add_inf_z, white)
                for k in ('rgb', 'depth', 'mask', 'weights', 'trans_shift', 'alpha'):
                    out[key + k] = r[k].numpy()

    for name in render_cases.DENSE_CASES:
        c = render_cases.dense_case(name)
        sigma, rad, z = (torch.from_numpy(c[k]) for k in ('sigma', 'rad', 'zvals'))
        bkg = torch.from_numpy(c['bkg']) if c['bkg'] is not None else None
        sigma.requires_grad_(True); rad.requires_grad_(True)
        r = ray_marching(sigma, rad, z, add_inf_z=c['add_inf_z'], white_bkg=c['white_bkg'], bkg_color=bkg)
        for k in ('rgb', 'depth', 'mask', 'weights', 'trans_shift', 'alpha'):
            out['dense_%s_%s' % (name, k)] = r[k].detach().numpy()
        # gradients of a fixed scalar functional of the outputs (autograd through the reference code)
        g_rgb, g_depth, g_mask = (torch.from_numpy(c[k]) for k in ('g_rgb', 'g_depth', 'g_mask'))
        loss = (r['rgb'] * g_rgb).sum() + (r['depth'] * g_depth).sum() + (r['mask'] * g_mask).sum()
        loss.backward()
        out['dense_%s_dsigma' % name] = sigma.grad.numpy()
        out['dense_%s_drad' % name] = rad.grad.numpy()

    # alpha input (NeuS style): ray_marching(sdf, radiance, mid_zvals, alpha=alpha)
    c = render_cases.alpha_case()
    alpha = torch.from_numpy(c['alpha']).requires_grad_(True)
    rad = torch.from_numpy(c['rad']).requires_grad_(True)
    r = ray_marching(torch.from_numpy(c['sdf']), rad, torch.from_numpy(c['zvals']), alpha=alpha,
                     bkg_color=torch.from_numpy(c['bkg']))
    for k in ('rgb', 'depth', 'mask', 'weights'):
        out['alpha_%s' % k] = r[k].detach().numpy()
    ((r['rgb'] * torch.from_numpy(c['g_rgb'])).sum() + (r['weights'] * torch.from_numpy(c['g_w'])).sum()).backward()
    out['alpha_dalpha'], out['alpha_drad'] = alpha.grad.numpy(), rad.grad.numpy()
    t, w = alpha_to_weights(torch.from_numpy(c['alpha']))
    out['alpha_a2w_trans'], out['alpha_a2w_weights'] = t.numpy(), w.numpy()

    # sdf_to_alpha + its gradients
    c = render_cases.sdf_case()
    sdf = torch.from_numpy(c['sdf']).requires_grad_(True)
    slope = torch.from_numpy(c['slope']).requires_grad_(True)
    s = torch.tensor(c['s'], requires_grad=True)
    a = sdf_to_alpha(sdf, torch.from_numpy(c['zvals']), slope, s)
    out['sdf_alpha'] = a.detach().numpy()
    (a * torch.from_numpy(c['g'])).sum().backward()
    out['sdf_dsdf'], out['sdf_dslope'], out['sdf_ds'] = sdf.grad.numpy(), slope.grad.numpy(), s.grad.numpy()

    # sample_pdf, deterministic u (linspace)
    c = render_cases.pdf_case()
    out['pdf_samples'] = sample_pdf(torch.from_numpy(c['bins']), torch.from_numpy(c['weights']), c['n_sample'], det=True).numpy()

    # SH basis: ArcNerf's own torch implementation evaluated on raw directions
    d = render_cases.unit_dirs(512, 3)
    for deg in (1, 2, 3, 4, 5):
        emb = SHEmbedder(n_freqs=deg, include_input=False, backend='torch', dtype='torch.float32')
        out['sh_torch_deg%d' % deg] = emb.sh_encode_torch(torch.from_numpy(d)).numpy()

    path = os.path.join(ROOT, 'tests', 'golden', 'render_golden.npz')
    np.savez_compressed(path, **out)
    print('wrote', path, os.path.getsize(path), 'bytes,', len(out), 'arrays')


if __name__ == '__main__':
    main()
