# -*- coding: utf-8 -*-
"""Model shells around the dense-chain networks: the reference's two-stage `NeRF` foreground model
(arcnerf/models/nerf_model.py:19-117 on top of FgModel.forward, fg_model.py:153-229, with the BasicBound that samples
near..far, basic_bound.py:46-106), its NeRF++ background model `NeRFPP` (nerfpp_bkg_model.py:20-136) and the
`FullModel` that chunks rays and blends the two (full_model.py:310-330, 415-475), with the same constructor fields,
forward arguments and output keys.  Everything a forward runs is a libarcb200 kernel: depth samples (csrc/zvals.cu), ray
points, both networks with their frequency embeddings (csrc/dense_chain.cu), compositing and inverse-CDF re-sampling
(csrc/composite.cu).  Gradients flow through the library's backward kernels (autograd.Functions), so the reference's
trainer, losses and optimiser can drive these modules unchanged.

Not mirrored: mask-based ray compaction (these two models sample every ray between near and far / on every shell, so
FgModel.forward takes its `mask_rays is None` branch), siren pre-training, surface rendering."""
import torch
import torch.nn as nn

from . import ray_helper as rh
from .nerf_mlp import FusedNeRFNets
from .registry import MODEL_REGISTRY, cfg_to_dict


def _get(d, key, default=None):
    v = d.get(key, default)
    return default if v is None else v


def _nets_from_cfgs(model_cfg, d_x, device):
    geo, rad = cfg_to_dict(model_cfg.get('geometry')), cfg_to_dict(model_cfg.get('radiance'))
    enc_g = cfg_to_dict(geo.get('encoder'))
    enc_v = cfg_to_dict(cfg_to_dict(rad.get('encoder')).get('view'))
    assert _get(enc_g, 'type', 'FreqEmbedder') == 'FreqEmbedder' and _get(enc_v, 'type', 'FreqEmbedder') == 'FreqEmbedder', \
        'the dense-chain networks take frequency embeddings (hash-grid models: NGPFgModel / NeusNGP)'
    assert _get(rad, 'mode', 'vf') == 'vf' and _get(rad, 'D', 1) == 1, "RadianceNet mode 'vf' with one hidden layer"
    assert _get(geo, 'W_feat', 256) == _get(geo, 'W', 256) == _get(rad, 'W_feat_in', 256), 'W_feat = W = W_feat_in'
    assert _get(enc_g, 'input_dim', d_x) == d_x
    return FusedNeRFNets(d_x, _get(enc_g, 'n_freqs', 10), _get(enc_v, 'n_freqs', 4), _get(geo, 'W', 256), _get(geo, 'D', 8),
                         _get(geo, 'skips', [4]), _get(rad, 'W', 128), device=device)


class _TwoStage(nn.Module):
    """what NeRF and NeRFPP share: coarse / fine network pairs, ray_marching with the ray cfgs, output naming"""

    def __init__(self, model_cfg, d_x, device):
        super().__init__()
        rays = cfg_to_dict(model_cfg.get('rays'))
        self.ray_cfgs = dict(
            bounding_radius=rays.get('bounding_radius'), near=rays.get('near'), far=rays.get('far'),
            n_sample=_get(rays, 'n_sample', 64), n_importance=_get(rays, 'n_importance', 0),
            inverse_linear=_get(rays, 'inverse_linear', False), perturb=_get(rays, 'perturb', False),
            add_inf_z=_get(rays, 'add_inf_z', False), noise_std=_get(rays, 'noise_std', 0.0), white_bkg=_get(rays, 'white_bkg', False),
            shared_network=_get(rays, 'shared_network', False))
        self.chunk_rays = _get(model_cfg, 'chunk_rays', 32768)
        self.coarse_nets = _nets_from_cfgs(model_cfg, d_x, device)
        if self.ray_cfgs['n_importance'] > 0:
            self.fine_nets = self.coarse_nets if self.ray_cfgs['shared_network'] else _nets_from_cfgs(model_cfg, d_x, device)

    def get_ray_cfgs(self, key=None):
        return self.ray_cfgs if key is None else self.ray_cfgs[key]

    def get_chunk_rays(self):
        return self.chunk_rays

    def get_net(self):
        return self.fine_nets if self.ray_cfgs['n_importance'] > 0 else self.coarse_nets

    def ray_marching(self, sigma, radiance, zvals, inference_only=False, bkg_color=None):
        """Base3dModel.ray_marching (base_3d_model.py:113-158)"""
        c = self.ray_cfgs
        return rh.ray_marching(sigma, radiance, zvals, c['add_inf_z'], c['noise_std'] if not inference_only else 0.0,
                               white_bkg=c['white_bkg'], bkg_color=bkg_color)

    @staticmethod
    def output_get_progress(output, get_progress):
        """base_3d_model.py:160-188: rgb / depth / mask always, the per-sample records as detached `progress_*` on request"""
        out = {k: output[k] for k in ('rgb', 'depth', 'mask')}
        if get_progress:
            for k in ('sigma', 'zvals', 'alpha', 'trans_shift', 'weights'):
                if output.get(k) is not None:
                    out['progress_' + k] = output[k].detach()
        return out

    def adjust_coarse_fine_output(self, output, inference_only):
        """base_3d_model.py:189-208"""
        if inference_only:
            return output['fine'] if self.ray_cfgs['n_importance'] > 0 else output['coarse']
        out = {k + '_coarse': v for k, v in output['coarse'].items()}
        if self.ray_cfgs['n_importance'] > 0:
            out.update({k + '_fine': v for k, v in output['fine'].items()})
        return out

    def _sigma_radiance(self, nets, pts, rays_d, n_pts):
        """points (B, n_pts, d) + one direction per ray -> sigma (B, n_pts), radiance (B, n_pts, 3)"""
        b = pts.shape[0]
        dirs = rays_d[:, None, :].expand(b, n_pts, 3).reshape(-1, 3)
        sigma, rgb = nets(pts.reshape(b * n_pts, -1), dirs)
        return sigma.view(b, n_pts), rgb.view(b, n_pts, 3)


@MODEL_REGISTRY.register()
class NeRF(_TwoStage):
    """cfgs: the `model` section of configs/models/nerf.yaml (rays / chunk_rays / geometry / radiance)"""

    def __init__(self, cfgs, device='cuda:0'):
        super().__init__(cfg_to_dict(cfgs), 3, device)

    def forward(self, inputs, inference_only=False, get_progress=False, cur_epoch=0, total_epoch=300000):
        rays_o, rays_d = inputs['rays_o'].float().contiguous(), inputs['rays_d'].float().contiguous()
        bkg_color = inputs.get('bkg_color')
        c = self.ray_cfgs
        with torch.no_grad():
            near, far = rh.get_near_far_from_rays(rays_o, rays_d, inputs.get('bounds'), c['near'], c['far'], c['bounding_radius'])
            zvals = rh.get_zvals_from_near_far(near, far, c['n_sample'], inverse_linear=c['inverse_linear'],
                                               perturb=c['perturb'] and not inference_only)
            pts = rh.get_ray_points_by_zvals(rays_o, rays_d, zvals)
        output = {}
        sigma, radiance = self._sigma_radiance(self.coarse_nets, pts, rays_d, c['n_sample'])
        out_c = self.ray_marching(sigma, radiance, zvals, inference_only, bkg_color)
        output['coarse'] = self.output_get_progress(out_c, get_progress)
        if c['n_importance'] > 0:
            with torch.no_grad():
                zvals, _ = rh.upsample_zvals(zvals, out_c['weights'].detach(), c['n_importance'],
                                             det=(not c['perturb']) if not inference_only else True)
                pts = rh.get_ray_points_by_zvals(rays_o, rays_d, zvals)
            sigma, radiance = self._sigma_radiance(self.fine_nets, pts, rays_d, c['n_sample'] + c['n_importance'])
            out_f = self.ray_marching(sigma, radiance, zvals, inference_only, bkg_color)
            output['fine'] = self.output_get_progress(out_f, get_progress)
        return self.adjust_coarse_fine_output(output, inference_only)

    @torch.no_grad()
    def forward_pts_dir(self, pts, view_dir=None):
        """Base3dModel.forward_pts_dir (base_3d_model.py:210-231): sigma (N,), rgb (N,3) at world points"""
        d = torch.zeros_like(pts) if view_dir is None else view_dir
        if view_dir is None:
            d = d.clone(); d[:, 2] = 1.0          # a zero direction has no normalisation; the reference feeds zeros through its norm-free path
        sigma, rgb = self.get_net()(pts.float().contiguous(), d.float().contiguous())
        return sigma[:, 0], rgb


@MODEL_REGISTRY.register()
class NeRFPP(_TwoStage):
    """NeRF++ background: inverse-sphere shells outside `bounding_radius`, points as (x/r, y/r, z/r, 1/r).
    cfgs: the `model.background` section of configs/models/nerfpp.yaml"""

    def __init__(self, cfgs, device='cuda:0'):
        super().__init__(cfg_to_dict(cfgs), 4, device)
        assert self.ray_cfgs['bounding_radius'] is not None, 'Please specify the bounding radius for nerf++ model'

    @staticmethod
    def _inverse_sphere_pts(rays_o, rays_d, zvals, radius=None):
        pts = rh.get_ray_points_by_zvals(rays_o, rays_d, zvals)
        if radius is None:
            radius = torch.norm(pts, dim=-1, keepdim=True)
        return torch.cat([pts / radius, 1.0 / radius.expand(pts.shape[0], pts.shape[1], 1)], -1)

    def forward(self, inputs, inference_only=False, get_progress=False, cur_epoch=0, total_epoch=300000):
        rays_o, rays_d = inputs['rays_o'].float().contiguous(), inputs['rays_d'].float().contiguous()
        c = self.ray_cfgs
        with torch.no_grad():
            zvals, radius = rh.get_zvals_outside_sphere(rays_o, rays_d, c['n_sample'], c['bounding_radius'],
                                                        perturb=c['perturb'] and not inference_only)
            pts = self._inverse_sphere_pts(rays_o, rays_d, zvals, radius[None, :, None])
        output = {}
        sigma, radiance = self._sigma_radiance(self.coarse_nets, pts, rays_d, c['n_sample'])
        out_c = self.ray_marching(sigma, radiance, zvals, inference_only)
        output['coarse'] = self.output_get_progress(out_c, get_progress)
        if c['n_importance'] > 0:
            with torch.no_grad():
                zvals, _ = rh.upsample_zvals(zvals, out_c['weights'].detach(), c['n_importance'],
                                             det=(not c['perturb']) if not inference_only else True)
                pts = self._inverse_sphere_pts(rays_o, rays_d, zvals)
            sigma, radiance = self._sigma_radiance(self.fine_nets, pts, rays_d, c['n_sample'] + c['n_importance'])
            out_f = self.ray_marching(sigma, radiance, zvals, inference_only)
            output['fine'] = self.output_get_progress(out_f, get_progress)
        return self.adjust_coarse_fine_output(output, inference_only)


class FullModel(nn.Module):
    """fg_model + optional bkg_model over flattened, chunked rays, blended as the reference's `bkg_blend: rgb` does
    (full_model.py:310-330,415-475): rgb = rgb_fg + T_fg_last * rgb_bkg, depth likewise, mask foreground only."""

    def __init__(self, fg_model, bkg_model=None):
        super().__init__()
        self.fg_model, self.bkg_model = fg_model, bkg_model

    def get_chunk_rays(self):
        c = self.fg_model.get_chunk_rays()
        return min(c, self.bkg_model.get_chunk_rays()) if self.bkg_model is not None else c

    @staticmethod
    def _blend(fg, bkg):
        """blend_bkg_rgb / blend_two_stage_bkg_rgb: each foreground stage takes the matching background stage (fine if present)"""
        two_stage_fg = any(k.endswith('_coarse') or k.endswith('_fine') for k in fg)
        b_rgb = bkg.get('rgb_fine', bkg.get('rgb_coarse', bkg.get('rgb')))
        b_depth = bkg.get('depth_fine', bkg.get('depth_coarse', bkg.get('depth')))
        out = dict(fg)
        for suffix in (('_coarse', '_fine') if two_stage_fg else ('',)):
            if 'rgb' + suffix not in fg:
                continue
            if two_stage_fg:
                br = bkg.get('rgb' + suffix, b_rgb)
                bd = bkg.get('depth' + suffix, b_depth)
            else:
                br, bd = b_rgb, b_depth
            lam = fg['progress_trans_shift' + suffix][:, -1]
            out['rgb' + suffix] = fg['rgb' + suffix] + lam[:, None] * br
            out['depth' + suffix] = fg['depth' + suffix] + lam * bd
        return out

    def process_fg_bkg_model(self, flat_inputs, inference_only, get_progress, cur_epoch, total_epoch):
        if self.bkg_model is not None and getattr(self.fg_model, 'blends_bkg_in_compositor', False):
            return self._fg_over_bkg_fused(flat_inputs, inference_only, cur_epoch, total_epoch)
        get_progress_fg = True if self.bkg_model is not None else get_progress
        fg = self.fg_model(flat_inputs, inference_only, get_progress_fg, cur_epoch, total_epoch)
        if self.bkg_model is None:
            return fg
        bkg = self.bkg_model(flat_inputs, inference_only, True, cur_epoch, total_epoch)
        out = self._blend(fg, bkg)
        if not get_progress:
            out = {k: v for k, v in out.items() if not k.startswith('progress_')}
        return out

    def _fg_over_bkg_fused(self, flat_inputs, inference_only, cur_epoch, total_epoch):
        """The packed NGP foreground (plugin/fg_model.py) keeps no per-sample progress tensors; the same blend
        rgb = rgb_fg + T_last * rgb_bkg (blend_bkg_rgb, full_model.py:310-330) is evaluated by its compositing kernel when the
        background model's colour is handed in as the per-ray background colour; T_last (the reference's trans_shift[:, -1], which
        leaves the last weighted interval out) comes back from the kernel for the depth blend.  The background model runs first;
        its gradient arrives through d rgb / d bkg = T_last.  The depth blend carries no gradient to the foreground through T_last
        (depth is not part of any loss of the reference's configurations)."""
        bkg = self.bkg_model(flat_inputs, inference_only, False, cur_epoch, total_epoch)
        b_rgb = bkg.get('rgb_fine', bkg.get('rgb_coarse', bkg.get('rgb')))
        b_depth = bkg.get('depth_fine', bkg.get('depth_coarse', bkg.get('depth')))
        fg_in = dict(flat_inputs, bkg_color=b_rgb)
        fg = self.fg_model(fg_in, inference_only, False, cur_epoch, total_epoch)
        out = dict(fg)
        for suffix in ('', '_coarse'):
            if 'depth' + suffix in fg:
                out['depth' + suffix] = fg['depth' + suffix] + self.fg_model.last_trans * b_depth
        return out

    def forward(self, inputs, inference_only=False, get_progress=False, cur_epoch=0, total_epoch=300000):
        """inputs: tensors of shape (B, N, ...) like the reference's feed_in; outputs (B, N, ...)"""
        b, n = inputs['rays_o'].shape[:2]
        flat = {k: (v.reshape(b * n, *v.shape[2:]) if torch.is_tensor(v) else v) for k, v in inputs.items()}
        chunk = self.get_chunk_rays()
        outs = []
        for s in range(0, b * n, chunk):
            part = {k: (v[s:s + chunk] if torch.is_tensor(v) else v) for k, v in flat.items()}
            outs.append(self.process_fg_bkg_model(part, inference_only, get_progress, cur_epoch, total_epoch))
        out = {k: (torch.cat([o[k] for o in outs], 0) if len(outs) > 1 else outs[0][k]) for k in outs[0]}
        return {k: v.reshape(b, n, *v.shape[1:]) for k, v in out.items()}
