# -*- coding: utf-8 -*-
"""Drop-in for the sampler / occupancy part of arcnerf/models/multivol_bkg_model.py:MultiVol: the
instant-ngp style background of nested volumes (n_cascade cubes, each twice the side of the previous one, the innermost
left to the foreground model unless `inclusive`).  Same method names and returns as the reference; the cascaded cone-step
march and the density-grid maintenance go to libarcb200 (plugin/multivol_func.py, plugin/bitfield_func.py), compositing
to plugin/ray_helper.py.

The networks are handed in (any modules with the reference's contracts: geo_net(pts) -> (sigma (N,1), feat),
radiance_net(pts, view_dirs, normals, feat) -> (N,3)), so the class works with the fused-MLP heads, the dense NeRF
networks or analytic stand-ins."""
import torch
import torch.nn as nn

from .bitfield_func import ema_grid_samples_nerf, splat_grid_samples
from .multivol_func import generate_grid_samples_multivol, sparse_sampling_in_multivol_bitfield, update_bitfield_multivol
from .ray_helper import get_ray_points_by_zvals, ray_marching
from .volume import Volume


class MultiVol(nn.Module):
    def __init__(self, geo_net, radiance_net, basic_volume, n_sample=1024, cone_angle=0.0, epoch_optim=16,
                 epoch_optim_warmup=256, ema_optim_decay=0.95, opa_thres=0.01, near_distance=0.0, chunk_pts=1 << 20):
        """basic_volume: dict(n_grid=128, side=..., origin=..., n_cascade=..., inclusive=False) as cfgs.model.basic_volume"""
        super(MultiVol, self).__init__()
        self.geo_net, self.radiance_net = geo_net, radiance_net
        vol = dict(basic_volume)
        self.n_cascade = int(vol.pop('n_cascade'))
        assert self.n_cascade > 1, 'You should have at least 2 cascades...'
        self.inclusive = bool(vol.pop('inclusive', False))
        vol.setdefault('n_grid', 128)
        self.n_grid = vol['n_grid']
        self.basic_volume = Volume(**vol)
        max_len = [x * 2 ** (self.n_cascade - 1) for x in self.basic_volume.get_len()]
        self.max_volume = Volume(origin=tuple(float(x) for x in self.basic_volume.get_origin()), xyz_len=max_len)
        self.n_sample = n_sample
        self.cone_angle = cone_angle
        self.min_step = self.basic_volume.get_diag_len() / n_sample
        self.max_step = self.max_volume.get_diag_len() / self.n_grid
        self.optim_cfgs = dict(epoch_optim=epoch_optim, epoch_optim_warmup=epoch_optim_warmup, ema_optim_decay=ema_optim_decay,
                               opa_thres=opa_thres, near_distance=near_distance)
        self.chunk_pts = chunk_pts
        self.n_elements = self.n_grid ** 3
        self.total_n_elements = self.n_elements * (self.n_cascade if self.inclusive else self.n_cascade - 1)
        self.register_buffer('density_bitfield', torch.full((self.total_n_elements // 8, ), 255, dtype=torch.uint8))
        self.register_buffer('density_grid', torch.zeros((self.total_n_elements, ), dtype=torch.float32))
        self.register_buffer('density_grid_tmp', torch.zeros((self.total_n_elements, ), dtype=torch.float32))
        self.ema_step = 0

    def get_optim_cfgs(self, key=None):
        return self.optim_cfgs if key is None else self.optim_cfgs[key]

    # ------------------------------------------------------------------ sampling (multivol_bkg_model.py:72-112)
    def get_near_far_from_rays(self, rays_o, rays_d):
        """slab test against the outermost volume with the cameras INSIDE it (near = 0, far = exit distance) -> (B,1) each;
        the reference uses its torch path here for the same reason (the CUDA op treats an inside origin as a miss)"""
        rng = self.max_volume.get_range().to(rays_o.device)
        inv = 1.0 / torch.where(rays_d.abs() < 1e-12, torch.full_like(rays_d, 1e-12), rays_d)
        t0, t1 = (rng[:, 0][None] - rays_o) * inv, (rng[:, 1][None] - rays_o) * inv
        near = torch.minimum(t0, t1).amax(dim=1, keepdim=True).clamp_min(0.0)
        far = torch.maximum(t0, t1).amin(dim=1, keepdim=True).clamp_min(0.0)
        far = torch.where(far < near, near, far)
        return near, far

    def get_zvals_from_near_far(self, near, far, n_pts, rays_o, rays_d, **kwargs):
        """-> zvals (B, n_pts), mask_pts (B, n_pts) bool through the multi-resolution bitfield"""
        return sparse_sampling_in_multivol_bitfield(
            rays_o, rays_d, near, far, n_pts, self.cone_angle, self.min_step, self.max_step, self.basic_volume.get_range(),
            self.max_volume.get_range(), self.n_grid, self.n_cascade, self.density_bitfield,
            near_distance=self.get_optim_cfgs('near_distance'), inclusive=self.inclusive)

    def _forward_pts_dir(self, pts, rays_d):
        sigma, feat = self.geo_net(pts)
        radiance = self.radiance_net(pts, rays_d, None, feat)
        return sigma[..., 0], radiance

    def get_sigma_radiance_by_mask_pts(self, rays_o, rays_d, zvals, mask_pts):
        """valid samples only through the networks, the padded tail repeats the ray's last valid sample (:158-200)"""
        n_rays, n_pts = zvals.shape
        pts = get_ray_points_by_zvals(rays_o, rays_d, zvals)
        dirs = rays_d.unsqueeze(1).expand(n_rays, n_pts, 3)
        pts_v, dirs_v = pts[mask_pts].view(-1, 3), dirs[mask_pts].reshape(-1, 3)
        sig, rad = [], []
        for a in range(0, pts_v.shape[0], self.chunk_pts):
            s, r = self._forward_pts_dir(pts_v[a:a + self.chunk_pts], dirs_v[a:a + self.chunk_pts])
            sig.append(s); rad.append(r)
        _sigma, _radiance = torch.cat(sig, 0), torch.cat(rad, 0)
        last = torch.cumsum(mask_pts.sum(dim=1), dim=0) - 1
        sigma = _sigma[last].unsqueeze(1).expand(n_rays, n_pts).clone()
        radiance = _radiance[last].unsqueeze(1).expand(n_rays, n_pts, 3).clone()
        sigma[mask_pts] = _sigma
        radiance[mask_pts] = _radiance
        return sigma, radiance

    def forward(self, inputs, inference_only=False, get_progress=False, cur_epoch=0, total_epoch=300000):
        """-> dict(rgb, depth, mask[, progress tensors as ray_marching returns them]) for the background (:114-148)"""
        rays_o, rays_d = inputs['rays_o'], inputs['rays_d']
        n_rays = rays_o.shape[0]
        near, far = self.get_near_far_from_rays(rays_o, rays_d)
        zvals, mask_pts = self.get_zvals_from_near_far(near, far, self.n_sample, rays_o, rays_d)
        max_num_pts = max(1, int(mask_pts.sum(dim=1).max()))
        zvals, mask_pts = zvals[:, :max_num_pts], mask_pts[:, :max_num_pts]
        valid = torch.any(mask_pts, dim=1)
        if not torch.all(valid):
            sigma = torch.zeros((n_rays, max_num_pts), dtype=torch.float32, device=rays_o.device)
            radiance = torch.zeros((n_rays, max_num_pts, 3), dtype=torch.float32, device=rays_o.device)
            if torch.any(valid):
                _s, _r = self.get_sigma_radiance_by_mask_pts(rays_o[valid], rays_d[valid], zvals[valid], mask_pts[valid])
                sigma[valid], radiance[valid] = _s, _r
        else:
            sigma, radiance = self.get_sigma_radiance_by_mask_pts(rays_o, rays_d, zvals, mask_pts)
        out = ray_marching(sigma, radiance, zvals, add_inf_z=False, noise_std=0.0)
        if get_progress:
            return {k if k in ('rgb', 'depth', 'mask') else 'progress_' + k: v for k, v in out.items()}
        return {k: out[k] for k in ('rgb', 'depth', 'mask')}

    # ------------------------------------------------------------------ occupancy (:150-261)
    def get_density_grid_mean(self):
        return float(self.density_grid.clamp_min(0.0).mean().item())

    @torch.no_grad()
    def get_est_opacity(self, dt, pts):
        """base_3d_model.py:368-389: density * dt"""
        out = []
        for a in range(0, pts.shape[0], self.chunk_pts):
            out.append(self.geo_net(pts[a:a + self.chunk_pts])[0][..., 0])
        return torch.cat(out, 0) * dt

    @torch.no_grad()
    def optimize(self, cur_epoch=0):
        epoch_optim, warm = self.get_optim_cfgs('epoch_optim'), self.get_optim_cfgs('epoch_optim_warmup')
        if cur_epoch <= 0 or epoch_optim is None or cur_epoch % epoch_optim != 0:
            return
        if warm is not None and cur_epoch < warm:
            self._update_density_grid(self.total_n_elements, 0, self.n_sample)
        else:
            self._update_density_grid(self.total_n_elements // 4, self.total_n_elements // 4, self.n_sample)

    def _update_density_grid(self, n_uniform, n_nonuniform, n_pts):
        rng = self.basic_volume.get_range()
        pos_u, idx_u = generate_grid_samples_multivol(self.density_grid, n_uniform, rng, self.ema_step, self.n_cascade,
                                                      self.n_grid, -0.01, self.inclusive)
        pos_n, idx_n = generate_grid_samples_multivol(self.density_grid, n_nonuniform, rng, self.ema_step, self.n_cascade,
                                                      self.n_grid, self.get_optim_cfgs('opa_thres'), self.inclusive)
        pos, idx = torch.cat([pos_u, pos_n], dim=0), torch.cat([idx_u, idx_n], dim=0)
        dt = self.basic_volume.get_diag_len() / float(n_pts)
        opacity = self.get_est_opacity(dt, pos)
        self.density_grid_tmp.zero_()
        self.density_grid_tmp = splat_grid_samples(opacity, idx, n_uniform + n_nonuniform, self.density_grid_tmp)
        self.density_grid = ema_grid_samples_nerf(self.density_grid_tmp, self.density_grid, self.total_n_elements,
                                                  self.get_optim_cfgs('ema_optim_decay'))
        self.density_bitfield = update_bitfield_multivol(self.density_grid, self.get_density_grid_mean(), self.density_bitfield,
                                                         self.get_optim_cfgs('opa_thres'), self.n_grid, self.n_cascade,
                                                         self.inclusive)
        self.ema_step = self.ema_step + 1
