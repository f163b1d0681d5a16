"""CPU oracle of the vanilla-NeRF foreground model, the NeRF++ background model and their blend.  TEST INFRASTRUCTURE ONLY:
used by tests/ and, timed, by bench.py's cpu_baseline / --impl reference legs of the `nerf` configuration.

torch-CPU restatement (fp32, torch.nn.functional.linear = the reference's nn.Linear arithmetic) of
  arcnerf/models/base_modules/encoding/freq_encoder.py:65-88                   FreqEmbedder
  arcnerf/models/base_modules/geo_rad_model/linear_network_module.py:174-197   GeoNet.forward (skip concat [h, embed])
  linear_network_module.py:317-335 + encoder_mlp_network.py:93-118            RadianceNet.forward, mode 'vf' = [view embed, feat]
  arcnerf/models/nerf_model.py:52-117                                          NeRF._forward / upsample_zvals
  arcnerf/models/nerfpp_bkg_model.py:51-136                                    NeRFPP.forward
  arcnerf/models/full_model.py:310-330                                         blend_bkg_rgb (two-stage foreground)
PARITY PINNING: tests/golden/models_golden.npz holds outputs of the reference's own FullModel (built from
configs/models/nerfpp.yaml and nerf.yaml by tests/golden/make_golden_models.py in the build container, weights from
tests/nerf_full_cases.py); tests/test_oracle_nerf.py checks this file against it.
"""
import torch
import torch.nn.functional as F

from . import render, zvals as oz


def freq_embed(x, n_freqs):
    out = [x]
    for k in range(n_freqs):
        out += [torch.sin(x * 2.0 ** k), torch.cos(x * 2.0 ** k)]
    return torch.cat(out, -1)


class Nets:
    """GeoNet + RadianceNet from the reference's state dicts (dict name -> tensor)"""

    def __init__(self, geo_state, rad_state, n_freq_x=10, n_freq_v=4, D=8, skips=(4,), requires_grad=False):
        t = lambda v: torch.as_tensor(v).clone().float().requires_grad_(requires_grad)
        self.geo = [(t(geo_state['layers.%d.weight' % i]), t(geo_state['layers.%d.bias' % i])) for i in range(D + 1)]
        self.rad = [(t(rad_state['layers.%d.weight' % i]), t(rad_state['layers.%d.bias' % i])) for i in range(2)]
        self.n_freq_x, self.n_freq_v, self.D, self.skips = n_freq_x, n_freq_v, D, tuple(skips)

    def parameters(self):
        return [p for wb in self.geo + self.rad for p in wb]

    def __call__(self, x, dirs):
        xe = freq_embed(x, self.n_freq_x)
        h = xe
        for i, (w, b) in enumerate(self.geo):
            h = F.linear(h, w, b)
            if i < self.D:
                h = torch.relu(h)
            if i in self.skips:
                h = torch.cat([h, xe], -1)
        sigma, feat = h[:, 0], h[:, 1:]
        ve = freq_embed(dirs / torch.norm(dirs, dim=-1, keepdim=True), self.n_freq_v)
        r = torch.relu(F.linear(torch.cat([ve, feat], -1), *self.rad[0]))
        return sigma, torch.sigmoid(F.linear(r, *self.rad[1]))


def _two_stage(coarse, fine, rays_o, rays_d, zvals, pts_fn, cfg, bkg_color, noise=None):
    """coarse pass -> inverse-CDF re-sampling (deterministic) -> fine pass; returns (out_coarse, out_fine) ray_marching dicts"""
    b, n = zvals.shape
    pts = pts_fn(zvals)
    d = rays_d[:, None, :].expand(b, n, 3).reshape(-1, 3)
    sigma, rad = coarse(pts.reshape(b * n, -1), d)
    out_c = render.ray_marching(sigma.view(b, n), rad.view(b, n, 3), zvals, cfg['add_inf_z'], None, cfg['white_bkg'], None, bkg_color)
    out_f = None
    if cfg['n_importance'] > 0:
        w = out_c['weights'].detach()[:, 1:n - 1]
        mid = 0.5 * (zvals[:, 1:] + zvals[:, :-1])
        u = torch.linspace(0.0, 1.0, cfg['n_importance'])[None].expand(b, cfg['n_importance']).contiguous()
        new = torch.sort(render.sample_pdf(mid, w, u), -1)[0]
        z2 = torch.sort(torch.cat([zvals, new], -1), -1)[0]
        n2 = z2.shape[1]
        pts = pts_fn(z2, fine_pass=True)
        d = rays_d[:, None, :].expand(b, n2, 3).reshape(-1, 3)
        sigma, rad = fine(pts.reshape(b * n2, -1), d)
        out_f = render.ray_marching(sigma.view(b, n2), rad.view(b, n2, 3), z2, cfg['add_inf_z'], None, cfg['white_bkg'], None, bkg_color)
    return out_c, out_f


def nerf_forward(coarse, fine, rays_o, rays_d, cfg, bkg_color=None):
    """NeRF foreground with a BasicBound: near / far hard-coded or from the bounding sphere, `n_sample` linear depths"""
    near, far = oz.get_near_far_from_rays(rays_o, rays_d, None, cfg.get('near'), cfg.get('far'), cfg.get('bounding_radius'))
    z = oz.get_zvals_from_near_far(near, far, cfg['n_sample'], inverse_linear=cfg.get('inverse_linear', False))
    pts_fn = lambda zz, fine_pass=False: rays_o[:, None, :] + rays_d[:, None, :] * zz[..., None]
    return _two_stage(coarse, fine, rays_o, rays_d, z, pts_fn, cfg, bkg_color)


def nerfpp_forward(coarse, fine, rays_o, rays_d, cfg):
    """NeRF++ background: shells of radius bounding_radius / t, t in (0,1); points (x/r, y/r, z/r, 1/r)"""
    z, radius = oz.get_zvals_outside_sphere(rays_o, rays_d, cfg['n_sample'], cfg['bounding_radius'])

    def pts_fn(zz, fine_pass=False):
        p = rays_o[:, None, :] + rays_d[:, None, :] * zz[..., None]
        r = torch.norm(p, dim=-1, keepdim=True) if fine_pass else radius[None, :, None].expand(p.shape[0], p.shape[1], 1)
        return torch.cat([p / r, 1.0 / r], -1)
    return _two_stage(coarse, fine, rays_o, rays_d, z, pts_fn, cfg, None)


def blend(fg_c, fg_f, bkg_c, bkg_f):
    """full_model.py:277-308 (`bkg_blend: rgb`, both models two-stage): every foreground stage + its last transmittance x the
    background of the same stage"""
    out = {}
    for tag, fg, b in (('coarse', fg_c, bkg_c), ('fine', fg_f, bkg_f if bkg_f is not None else bkg_c)):
        if fg is None:
            continue
        lam = fg['trans_shift'][:, -1]
        out['rgb_' + tag] = fg['rgb'] + lam[:, None] * b['rgb']
        out['depth_' + tag] = fg['depth'] + lam * b['depth']
        out['mask_' + tag] = fg['mask']
    return out
