# -*- coding: utf-8 -*-
"""NeuS foreground model on libarcb200 (BASELINE.json configs[2], the NeuS-NGP configuration
configs/expr/NeRF/lego/nerf_lego_neus_ngp.yaml): the reference's `Neus(SdfModel(FgModel))` -
arcnerf/models/neus_model.py:19-265, sdf_model.py:42-114, fg_model.py:153-387 - with the same cfg fields, forward
arguments and output keys (rgb, depth, mask, normal, normal_pts, params, progress_*).

Per forward: occupancy-volume sampling (VolumeBound -> arcb200 ray-AABB + march kernels), `handle_mid_pts`, the SDF network
with normals (`geo_net.forward_with_grad`: hash grid with dy/dx + the 64-wide weight-normed softplus net on arcb200_gemm_nt),
the 'pvnf' radiance head (tcgen05 fused MLP, 38 inputs padded to 48 with ones), slope and cosine annealing, `sdf_to_alpha`
with the learnable scale s = exp(10 inv_s) (csrc/composite.cu), compositing from alpha, normal map.  The training backward
runs through the library's backward kernels, including the second-order pass the eikonal loss on `normal_pts` needs
(arcb200_grid_bwd_bwd_input, the twice-differentiable small-GEMM functions).

What stays in eager torch, as in the reference: the boolean-mask bookkeeping between padded (rays, points) matrices and the
list of valid points (fg_model.py:188-229, sdf_model.py:61-101), and elementwise glue (slope dot product, annealing, normal
normalisation, weight-norm re-parametrisation of 5 k parameters)."""
import math

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as tnf

from . import ray_helper as rh
from .linear_network_module import GeoNet
from .obj_bound import VolumeBound, build_obj_bound
from .registry import MODEL_REGISTRY, MODULE_REGISTRY, cfg_to_dict
from .tcnn_fusedmlp_module import FusedMLPRadianceNet  # noqa: F401  (registers the module)


def _get(d, key, default=None):
    v = d.get(key, default)
    return default if v is None else v


def _build(cfg, default_type):
    d = cfg_to_dict(cfg)
    cls = MODULE_REGISTRY.get(d.get('type') or default_type)
    return cls(**{k: v for k, v in d.items() if k != 'type'})


@MODEL_REGISTRY.register()
class Neus(nn.Module):
    """cfgs: the `model` section of nerf_lego_neus_ngp.yaml (obj_bound / rays / params / chunk_* / geometry / radiance)"""

    def __init__(self, cfgs, device='cuda:0'):
        super().__init__()
        c = cfg_to_dict(cfgs)
        rays, params, bound = cfg_to_dict(c.get('rays')), cfg_to_dict(c.get('params')), cfg_to_dict(c.get('obj_bound'))
        self.ray_cfgs = dict(n_sample=_get(rays, 'n_sample', 128), n_importance=_get(rays, 'n_importance', 0), n_iter=_get(rays, 'n_iter', 4),
                             inverse_linear=_get(rays, 'inverse_linear', False), perturb=_get(rays, 'perturb', False),
                             add_inf_z=_get(rays, 'add_inf_z', False), noise_std=_get(rays, 'noise_std', 0.0),
                             white_bkg=_get(rays, 'white_bkg', False), radius_bound=_get(rays, 'radius_bound', 1.5))
        assert not self.ray_cfgs['add_inf_z'], 'Do not add_inf_z for sdf modeling'          # neus_model.py: the sdf models assert the same
        self.render_cfgs = dict(bkg_color=_get(bound, 'bkg_color', [1.0, 1.0, 1.0]), depth_far=_get(bound, 'depth_far', 10.0),
                                normal=_get(bound, 'normal', [0.0, 1.0, 0.0]))
        self.chunk_rays, self.chunk_pts = _get(c, 'chunk_rays', 32768), _get(c, 'chunk_pts', 1 << 20)
        self.geo_net = _build(c.get('geometry'), 'GeoNet')
        self.radiance_net = _build(c.get('radiance'), 'RadianceNet')
        self.speed_factor = _get(params, 'speed_factor', 10)
        self.inv_s = nn.Parameter(torch.tensor([-np.log(_get(params, 'init_var', 0.05)) / self.speed_factor], dtype=torch.float32))
        self.anneal_end = _get(params, 'anneal_end', 0)
        self.obj_bound, _ = build_obj_bound(c) if c.get('obj_bound') is not None else (VolumeBound(bound), 'volume')
        self.to(device)

    # ------------------------------------------------------------------------------------------ small accessors
    def get_ray_cfgs(self, key=None):
        return self.ray_cfgs if key is None else self.ray_cfgs[key]

    def get_chunk_rays(self):
        return self.chunk_rays

    def forward_scale(self):
        """scale = exp(inv_s * speed)   (neus_model.py:52-54)"""
        return torch.exp(self.inv_s * self.speed_factor)

    def get_cos_anneal(self, cur_epoch):
        return 1.0 if self.anneal_end == 0 else min(1.0, cur_epoch / self.anneal_end)

    # ------------------------------------------------------------------------------------------ FgModel.forward
    def forward(self, inputs, inference_only=False, get_progress=False, cur_epoch=0, total_epoch=300000):
        rays_o, rays_d = inputs['rays_o'].float().contiguous(), inputs['rays_d'].float().contiguous()
        bkg_color = inputs.get('bkg_color')
        n_rays = rays_o.shape[0]
        with torch.no_grad():
            near, far, mask_rays = self.obj_bound.get_near_far_from_rays({'rays_o': rays_o, 'rays_d': rays_d})
            zvals, mask_pts = self.obj_bound.get_zvals_from_near_far(near, far, self.ray_cfgs['n_sample'], inference_only,
                                                                     self.ray_cfgs['inverse_linear'], self.ray_cfgs['perturb'],
                                                                     rays_o=rays_o, rays_d=rays_d)
            if mask_pts is None:                          # bounds without an occupancy structure (sphere / near-far): every sample counts
                mask_pts = torch.ones(zvals.shape, dtype=torch.bool, device=zvals.device)
            if mask_rays is None:
                mask_rays = torch.ones(n_rays, dtype=torch.bool, device=zvals.device)
            # the three data-dependent decisions of FgModel.forward (longest ray, any / all rays valid) in ONE host read
            per_ray = mask_pts.sum(dim=1)
            valid0 = torch.logical_and(mask_rays, per_ray > 0)
            stats = torch.stack([per_ray.max(), valid0.sum()]).tolist()
            max_num_pts = max(2, int(stats[0]))                                        # reduce_empty_mask (fg_model.py:251-262)
            zvals, mask_pts = zvals[:, :max_num_pts].contiguous(), mask_pts[:, :max_num_pts].contiguous()
            mask_rays, n_valid = valid0, int(stats[1])
        if n_valid == n_rays:
            return self._forward(rays_o, rays_d, zvals, mask_pts, bkg_color, inference_only, get_progress, cur_epoch)
        empty_batch = n_valid == 0
        if empty_batch:                                   # all rays miss: one synthetic ray gives the output keys (fg_model.py:198-207)
            ray_idx = torch.zeros(1, dtype=torch.long, device=zvals.device)
            z_valid = torch.zeros((1, zvals.shape[1]), dtype=zvals.dtype, device=zvals.device)
            z_valid[0, 1:] = 1.0
            m_valid = torch.zeros((1, mask_pts.shape[1]), dtype=torch.bool, device=zvals.device)
            m_valid[0, :2] = True
        else:
            # row indices of the valid rays, once; every per-ray tensor is gathered / scattered with them (a boolean mask would
            # run its own compaction - and a host synchronisation - at every use)
            ray_idx = torch.nonzero(mask_rays).squeeze(1)
            z_valid, m_valid = zvals.index_select(0, ray_idx), mask_pts.index_select(0, ray_idx)
        out_valid = self._forward(rays_o.index_select(0, ray_idx), rays_d.index_select(0, ray_idx), z_valid, m_valid,
                                  bkg_color.index_select(0, ray_idx) if bkg_color is not None else None, inference_only, get_progress,
                                  cur_epoch)
        return self.update_values_for_invalid_rays(out_valid, None if empty_batch else ray_idx, bkg_color, n_rays)

    def update_values_for_invalid_rays(self, out_valid, ray_idx, rand_bkg_color, n_rays):
        """fg_model.py:320-387: background colour / far depth / zero mask / default normal for the rays that miss the bound.
        ray_idx: rows of the valid rays (None: no ray was valid, out_valid only supplies keys and shapes)"""
        out = {}
        for k, v in out_valid.items():
            if not torch.is_tensor(v):
                out[k] = v
                continue
            shape = (n_rays, *v.shape[1:])
            kw = dict(dtype=v.dtype, device=v.device)
            if k.startswith('rgb'):
                fill = rand_bkg_color if rand_bkg_color is not None else torch.tensor(self.render_cfgs['bkg_color'], **kw)[None]
                t = torch.ones(shape, **kw) * fill
            elif k.startswith('depth'):
                t = torch.ones(shape, **kw) * self.render_cfgs['depth_far']
            elif k.startswith('mask'):
                t = torch.zeros(shape, **kw)
            elif k.startswith('normal'):
                nrm = tnf.normalize(torch.tensor(self.render_cfgs['normal'], **kw)[None], dim=-1)
                t = torch.ones(shape, **kw) * (nrm[None] if k == 'normal_pts' else nrm)
            elif k.startswith('progress'):
                t = -torch.ones(shape, **kw) if 'sigma' in k else (torch.ones(shape, **kw) if 'trans_shift' in k else torch.zeros(shape, **kw))
            else:
                out[k] = v
                continue
            if ray_idx is not None:
                t = t.index_copy(0, ray_idx, v)
            out[k] = t
        return out

    # ------------------------------------------------------------------------------------------ Neus._forward
    def handle_mid_pts(self, zvals, mask_pts):
        """neus_model.py:173-202 (the masked branch): interval mid points, one extra interval of the ray's step appended"""
        n_rays, n_pts = zvals.shape
        sample_dist = (zvals[:, -1] - zvals[:, 0]) / self.ray_cfgs['n_sample'] * 0.5
        final_zvals = zvals[:, -1] + sample_dist * 2.0
        _zvals = (torch.ones((n_rays, n_pts + 1), dtype=zvals.dtype, device=zvals.device) * final_zvals.unsqueeze(1)).contiguous()
        # _zvals[_mask] = zvals[mask_pts] with _mask = [mask_pts | False]: the same entries, one column wider
        idx = torch.nonzero(mask_pts.reshape(-1)).squeeze(1)
        ray = torch.div(idx, n_pts, rounding_mode='floor')
        _zvals.view(-1).index_copy_(0, idx + ray, zvals.reshape(-1).index_select(0, idx))
        mid_zvals = 0.5 * (_zvals[..., 1:] + _zvals[..., :-1])
        mask_mid = torch.cat([torch.ones_like(mask_pts[:, :1]), mask_pts[:, :-1]], dim=1)
        return mid_zvals, _zvals, mask_mid

    def get_sdf_radiance_normal_by_mask_pts(self, rays_o, rays_d, zvals, mask_pts):
        """sdf_model.py:42-114: networks on the valid points only, results re-expanded with the ray's last valid value.
        The valid entries are addressed by ONE index tensor (flat position, ray) instead of a boolean mask per tensor."""
        n_rays, n_pts = zvals.shape
        pts = rh.get_ray_points_by_zvals(rays_o, rays_d, zvals)
        idx = torch.nonzero(mask_pts.reshape(-1)).squeeze(1)
        ray = torch.div(idx, n_pts, rounding_mode='floor')
        pts_v, dirs_v = pts.view(-1, 3).index_select(0, idx), rays_d.index_select(0, ray)
        sdf_v, rad_v, nrm_v = self._forward_pts_dir(pts_v, dirs_v)
        last = torch.cumsum(mask_pts.sum(dim=1), dim=0) - 1
        fill = lambda v, c: (v.index_select(0, last).unsqueeze(1).expand(n_rays, n_pts, *v.shape[1:]).reshape(n_rays * n_pts, *v.shape[1:]))
        sdf = fill(sdf_v, 1).index_copy(0, idx, sdf_v).view(n_rays, n_pts)
        rad = fill(rad_v, 3).index_copy(0, idx, rad_v).view(n_rays, n_pts, 3)
        nrm = fill(nrm_v, 3).index_copy(0, idx, nrm_v).view(n_rays, n_pts, 3)
        return sdf, rad, nrm

    def _forward_pts_dir(self, pts, rays_d):
        """sdf_model.py:104-114: sdf (N,), radiance (N,3), normal (N,3) = d sdf / d pts kept in the graph"""
        sdf, feat, normal = self.geo_net.forward_with_grad(pts)
        radiance = self.radiance_net(pts, rays_d, normal, feat)
        return sdf[..., 0], radiance, normal

    def surface_render(self, inputs, method='sphere_tracing', n_step=128, n_iter=20, threshold=0.01, level=0.0, grad_dir='ascent'):
        """SdfModel.surface_render (sdf_model.py:116-171): the first surface point of every ray by sphere tracing or the secant
        method on the SDF network (plugin/surface.py: no host synchronisation inside the search), then colour and normal from the
        networks at that point.  Returns rgb (B,3) (white where nothing is hit), depth (B,1), mask (B,), normal (B,3)."""
        from .surface import surface_ray_intersection
        assert level == 0.0, 'Invalid level for sdf model...'
        assert grad_dir == 'ascent', 'Invalid grad_dir for sdf model...'
        rays_o, rays_d = inputs['rays_o'].float().contiguous(), inputs['rays_d'].float().contiguous()
        n_rays, dev = rays_o.shape[0], rays_o.device
        with torch.no_grad():
            near, far, valid_rays = self.obj_bound.get_near_far_from_rays({'rays_o': rays_o, 'rays_d': rays_d})
        geo_value = self.geo_net.forward_geo_value
        if valid_rays is None or bool(torch.all(valid_rays)):
            zvals, pts, mask = surface_ray_intersection(rays_o, rays_d, geo_value, method, near, far, n_step, n_iter, threshold, level, grad_dir)
        else:
            zv, pv, mv = surface_ray_intersection(rays_o[valid_rays], rays_d[valid_rays], geo_value, method, near[valid_rays],
                                                  far[valid_rays], n_step, n_iter, threshold, level, grad_dir)
            zvals = torch.ones((n_rays, 1), dtype=torch.float32, device=dev) * (zv.max() if zv.numel() else 0.0)
            pts = torch.ones((n_rays, 3), dtype=torch.float32, device=dev)
            mask = torch.zeros((n_rays, ), dtype=torch.bool, device=dev)
            zvals[valid_rays], pts[valid_rays], mask[valid_rays] = zv, pv, mv
        rgb = torch.ones((n_rays, 3), dtype=torch.float32, device=dev)                  # white bkg
        normal = torch.zeros((n_rays, 3), dtype=torch.float32, device=dev)
        if bool(torch.any(mask)):
            _, rgb_m, normal_m = self._forward_pts_dir(pts[mask], rays_d[mask])
            rgb[mask], normal[mask] = rgb_m.detach().float(), normal_m.detach().float()
        return {'rgb': rgb, 'depth': zvals, 'mask': mask.float(), 'normal': normal}

    @torch.no_grad()
    def forward_pts(self, pts):
        """sdf at world points (N, 3) -> (N,) in chunks of chunk_pts (SdfModel.forward_pts, sdf_model.py:19-40)"""
        out = [self.geo_net.forward_geo_value(pts[a:a + self.chunk_pts]) for a in range(0, pts.shape[0], self.chunk_pts)]
        return out[0] if len(out) == 1 else torch.cat(out, 0)

    @torch.no_grad()
    def upsample_zvals(self, rays_o, rays_d, zvals, mask_pts=None, inference_only=False, s=32):
        """neus_model.py:106-171: n_iter rounds of inverse-CDF re-sampling where the SDF drops fastest along the ray (eq. 13 of the
        NeuS paper); the SDF network, sdf_to_alpha, the inverse-CDF gather and the row sort run on the library's kernels"""
        c = self.ray_cfgs
        if c['n_importance'] <= 0:
            return zvals, mask_pts
        n_per_iter = c['n_importance'] // c['n_iter']
        det = (not c['perturb']) if not inference_only else True
        for i in range(c['n_iter']):
            n_rays, n_pts = zvals.shape
            pts = rh.get_ray_points_by_zvals(rays_o, rays_d, zvals)                    # (B, N, 3)
            sdf = self.forward_pts(pts.view(-1, 3)).view(n_rays, n_pts)
            prev_sdf, next_sdf = sdf[:, :-1], sdf[:, 1:]
            mid_sdf = (prev_sdf + next_sdf) * 0.5
            slope = (next_sdf - prev_sdf) / (zvals[:, 1:] - zvals[:, :-1] + 1e-5)
            prev_slope = torch.cat([torch.zeros((n_rays, 1), dtype=zvals.dtype, device=zvals.device), slope[:, :-1]], dim=-1)
            slope = torch.minimum(prev_slope, slope).clamp(-10.0, 0.0)
            radius = torch.norm(pts, dim=-1)
            inside = (radius[:, :-1] < c['radius_bound']) | (radius[:, 1:] < c['radius_bound'])
            slope = slope * inside
            alpha = rh.sdf_to_alpha(mid_sdf.contiguous(), zvals.contiguous(), slope.contiguous(), float(s * (2 ** (i + 1))), clip=False)
            _, weights = rh.alpha_to_weights(alpha)
            new = rh.sample_pdf(zvals, weights, n_per_iter, det).detach()
            zvals = rh.sort_rows(torch.cat([zvals, new], dim=-1))
            if mask_pts is not None:                      # merge_full_mask (fg_model.py:390-406): the new samples are all valid
                count = mask_pts.sum(dim=1, keepdim=True) + new.shape[1]
                mask_pts = torch.arange(zvals.shape[1], device=zvals.device)[None, :] < count
        return zvals, mask_pts

    def _forward(self, rays_o, rays_d, zvals, mask_pts, bkg_color, inference_only, get_progress, cur_epoch):
        zvals, mask_pts = self.upsample_zvals(rays_o, rays_d, zvals, mask_pts, inference_only)
        mid_zvals, zvals, mask_mid = self.handle_mid_pts(zvals, mask_pts)
        sdf, radiance, normal_pts = self.get_sdf_radiance_normal_by_mask_pts(rays_o, rays_d, mid_zvals, mask_mid)
        cos_anneal = 1.0 if inference_only else self.get_cos_anneal(cur_epoch)
        slope = torch.sum(rays_d[:, None, :] * normal_pts, dim=-1)
        iter_slope = -(tnf.relu(-slope * 0.5 + 0.5) * (1 - cos_anneal) + tnf.relu(-slope) * cos_anneal)
        alpha = rh.sdf_to_alpha(sdf.contiguous(), zvals.contiguous(), iter_slope.contiguous(), self.forward_scale())
        out = rh.ray_marching(sdf, radiance.contiguous(), mid_zvals.contiguous(), False, 0.0, white_bkg=self.ray_cfgs['white_bkg'],
                              alpha=alpha, bkg_color=bkg_color)
        output = {k: out[k] for k in ('rgb', 'depth', 'mask')}
        output['normal'] = torch.sum(out['weights'].unsqueeze(-1) * tnf.normalize(normal_pts, dim=-1), -2)
        if not inference_only:
            output['params'] = {'scale': float(self.forward_scale().detach())}
            output['normal_pts'] = normal_pts
        if get_progress:
            for k in ('sigma', 'zvals', 'alpha', 'trans_shift', 'weights'):
                if out.get(k) is not None:
                    output['progress_' + k] = out[k].detach()
        return output

    # ------------------------------------------------------------------------------------------ occupancy refresh
    @torch.no_grad()
    def _geo_with_normal(self, pts):
        with torch.enable_grad():
            sdf, _, normal = self.geo_net.forward_with_grad(pts.clone())
        return sdf.detach(), normal.detach()

    def get_est_opacity(self, dt, pts):
        """neus_model.py:204-219: opacity of a fake ray pointing at the origin through every sampled voxel point"""
        n = pts.shape[0]
        rays_d = -tnf.normalize(pts, dim=-1)
        sdf, normal = self._geo_with_normal(pts)
        slope = torch.sum(rays_d * normal, dim=-1, keepdim=True)
        zvals = torch.zeros((n, 2), dtype=pts.dtype, device=pts.device)
        zvals[:, 1] += dt * 1.0 / math.sqrt(3)
        with torch.no_grad():
            return rh.sdf_to_alpha(sdf.contiguous(), zvals, (-tnf.relu(-slope)).contiguous(), self.forward_scale().detach())[:, 0]

    def optimize(self, cur_epoch=0):
        """FgModel.optimize (fg_model.py:400-410): refresh the occupancy volume from the current SDF"""
        self.obj_bound.optimize(cur_epoch, self.ray_cfgs['n_sample'], self.get_est_opacity)


def eikonal_loss(normal_pts):
    """EikonalLoss (arcnerf/loss/geo_loss.py:12-73, key normal_pts): mean (|grad sdf| - 1)^2 over all points"""
    return ((torch.norm(normal_pts, dim=-1) - 1.0) ** 2).mean()
