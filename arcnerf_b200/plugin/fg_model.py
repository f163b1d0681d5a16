# -*- coding: utf-8 -*-
"""Foreground model of the NGP configuration on the fused pipeline: the reference's
`NeRF(FgModel)` with an occupancy-volume bound, hash-grid + fused-MLP heads and n_importance = 0
(arcnerf/models/fg_model.py:153-410, nerf_model.py:52-91, full_model.py:310-330,415-475), as ONE module with the same
`forward(inputs, inference_only, get_progress, cur_epoch, total_epoch)` -> dict contract.

What the reference does with boolean masks, `nonzero`, dense re-expansion and per-chunk `torch.cat`
(`reduce_empty_mask`, `get_sigma_radiance_by_mask_pts`, `update_values_for_invalid_rays`, `chunk_processing`) happens
here inside the packed sample layout of arcnerf_b200.ngp.NGPPipeline: samples are emitted packed by the march, rays that
miss the bound simply own zero samples, and their defaults (background colour, depth_far, mask 0) are filled in at the end.

The forward is differentiable: parameters are `nn.Parameter`s that alias the pipeline's fp32 masters, gradients come from
the library's backward kernels through one autograd.Function, so the reference's trainer (its losses, torch optimiser,
DDP) can drive this module unchanged.  `train_step` is the fused alternative (loss + Adam inside the library).
"""
import torch
import torch.nn as nn

from ..ngp import NGP_DEFAULTS, NGPPipeline, init_params


class _RenderFn(torch.autograd.Function):
    """rays -> (rgb, depth, mask) through the pipeline's kernels; backward fills the pipeline's flat gradient buffer and
    hands its three views back as the gradients of the parameter tensors"""

    @staticmethod
    def forward(ctx, model, rays_o, rays_d, bkg, table, geo, rad):
        pipe = model.pipe
        model._sync_if_changed()
        R = rays_o.shape[0]
        rgb, depth, mask = pipe.forward(rays_o, rays_d, bkg)
        ctx.model, ctx.R, ctx.bkg, ctx.rays_d = model, R, bkg, rays_d
        ctx.set_key = pipe.forward_serial            # counts every sample() of the pipeline: two alternating sets would alias
        t_last = pipe.t_last[:R].clone()
        # a background MODEL behind this foreground (FullModel: bkg is its colour per ray) receives d rgb / d bkg = T_last
        ctx.trans_last = t_last if bkg.requires_grad else None
        ctx.mark_non_differentiable(t_last)
        return rgb.clone(), depth.clone(), mask.clone(), pipe.counts[:R].clone(), t_last

    @staticmethod
    def backward(ctx, g_rgb, g_depth, g_mask, _g_counts, _g_t_last):
        pipe, R = ctx.model.pipe, ctx.R
        if ctx.set_key != pipe.forward_serial:
            raise RuntimeError('backward through a chunk whose samples were overwritten by a later forward: '
                               'differentiate one chunk at a time (training batches fit one chunk)')
        pipe.g_all.zero_()
        if g_rgb is None:
            pipe.g_rgb[:R].zero_()
        else:
            pipe.g_rgb[:R].copy_(g_rgb)
        gd = g_depth.contiguous().float() if g_depth is not None else None
        gm = g_mask.contiguous().float() if g_mask is not None else None
        pipe.backward(ctx.rays_d, R, ctx.bkg, composite=True, g_depth=gd, g_mask=gm)
        g_bkg = None
        if ctx.trans_last is not None and g_rgb is not None:
            g_bkg = g_rgb * ctx.trans_last[:, None]
            if ctx.bkg.shape[0] == 1:
                g_bkg = g_bkg.sum(0, keepdim=True)
        return None, None, None, g_bkg, pipe.g_table.clone(), pipe.g_geo.clone(), pipe.g_rad.clone()


class NGPFgModel(nn.Module):
    """Drop-in for the reference's foreground model in the NGP configuration.

    cfgs (dict, optional): `rays` {n_sample 1024, near_distance ...}, `obj_bound` {volume {n_grid, side}, epoch_optim,
    epoch_optim_warmup, near_distance, opa_thres}, `render` {bkg_color [1,1,1], depth_far 10.0}, `chunk_rays` 32768 -
    the fields of configs/expr/NeRF/lego/nerf_lego_nerf_ngp.yaml that shape this path; everything else of NGP_DEFAULTS can
    be overridden through **overrides.
    """

    def __init__(self, cfgs=None, device='cuda:0', seed=1337, params=None, max_rays=None, capacity=None, **overrides):
        super(NGPFgModel, self).__init__()
        cfgs = cfgs or {}
        rays = cfgs.get('rays', {})
        bound = cfgs.get('obj_bound', {})
        render = cfgs.get('render', {})
        over = dict(overrides)
        if 'n_sample' in rays:
            over['n_sample'] = rays['n_sample']
        if 'volume' in bound:
            over['n_grid'] = bound['volume'].get('n_grid', 128)
            over['side'] = bound['volume'].get('side', NGP_DEFAULTS['side'])
        for k in ('near_distance', 'epoch_optim', 'epoch_optim_warmup', 'opa_thres'):
            if k in bound:
                over[k] = bound[k]
        self.chunk_rays = int(cfgs.get('chunk_rays', 32768))
        self.bkg_color = [float(x) for x in render.get('bkg_color', bound.get('bkg_color', [1.0, 1.0, 1.0]))]
        self.depth_far = float(render.get('depth_far', 10.0))
        max_rays = int(max_rays or self.chunk_rays)
        capacity = int(capacity or max(1 << 19, max_rays * 48))
        cfg = dict(NGP_DEFAULTS)
        cfg.update(over)
        if params is None:
            params = init_params(cfg, seed)
        self.pipe = NGPPipeline(max_rays=max_rays, capacity=capacity, device=device, params=params, **over)
        # the pipeline's fp32 masters ARE the parameters (no copies): an external optimiser updates them in place
        self.table = nn.Parameter(self.pipe.table)
        self.geo_params = nn.Parameter(self.pipe.geo_p)
        self.rad_params = nn.Parameter(self.pipe.rad_p)
        self._versions = None
        self._remember_versions()

    # ------------------------------------------------------------------ parameter bookkeeping
    def _remember_versions(self):
        self._versions = (self.table._version, self.geo_params._version, self.rad_params._version)

    def _sync_if_changed(self):
        """fp16 table / packed MLP operands follow the fp32 parameters whenever something wrote to them"""
        now = (self.table._version, self.geo_params._version, self.rad_params._version)
        if now != self._versions:
            self.pipe.sync_params()
            self._versions = now

    # ------------------------------------------------------------------ the reference's model surface
    blends_bkg_in_compositor = True      # FullModel: a background model's colour goes in as the per-ray bkg_color

    def get_chunk_rays(self):
        return min(self.chunk_rays, self.pipe.max_rays)

    def get_n_coarse_sample(self):
        return self.pipe.cfg['n_sample']

    def get_obj_bound(self):
        return getattr(self.pipe, 'bound', None)

    @torch.no_grad()
    def forward_pts(self, pts):
        """density at world points (N,3) -> (N,)   (base_3d_model.py:343-366)"""
        self._sync_if_changed()
        return self.pipe.forward_pts(pts)

    @torch.no_grad()
    def get_est_opacity(self, dt, pts):
        self._sync_if_changed()
        return self.pipe.get_est_opacity(dt, pts)

    @torch.no_grad()
    def optimize(self, cur_epoch=0):
        """FgModel.optimize (fg_model.py:400-410): refresh the occupancy every `epoch_optim` steps"""
        self._sync_if_changed()
        self.pipe.update_occupancy(cur_epoch)

    def forward(self, inputs, inference_only=False, get_progress=False, cur_epoch=0, total_epoch=300000):
        """inputs: rays_o, rays_d (B,3), bkg_color (B,3) | (1,3) | None.  Returns rgb (B,3), depth (B,), mask (B,) under the
        keys `rgb/depth/mask` (inference_only) or `rgb_coarse/depth_coarse/mask_coarse` (training), as
        adjust_coarse_fine_output does for n_importance = 0 (base_3d_model.py:189-208).  Rays without a sample get the
        background colour, depth_far and mask 0 (update_values_for_invalid_rays, fg_model.py:318-387)."""
        if get_progress:
            raise NotImplementedError('per-sample progress tensors are not materialised by the packed pipeline')
        rays_o = inputs['rays_o'].to(self.pipe.dev, torch.float32).contiguous()
        rays_d = inputs['rays_d'].to(self.pipe.dev, torch.float32).contiguous()
        bkg = inputs.get('bkg_color', None)
        if bkg is None:
            bkg = torch.tensor([self.bkg_color], dtype=torch.float32, device=self.pipe.dev)
        bkg = bkg.to(self.pipe.dev, torch.float32)
        bkg = bkg if bkg.is_contiguous() else bkg.contiguous()
        n, chunk = rays_o.shape[0], self.pipe.max_rays
        grad = torch.is_grad_enabled() and not inference_only
        was_training, self.pipe.training = self.pipe.training, bool(grad)
        try:
            outs = []
            for s in range(0, n, chunk):
                e = min(n, s + chunk)
                b = bkg if bkg.shape[0] == 1 else bkg[s:e]
                if grad:
                    if n > chunk:
                        raise RuntimeError('a differentiable forward takes at most %d rays (one chunk); got %d' % (chunk, n))
                    outs.append(_RenderFn.apply(self, rays_o[s:e], rays_d[s:e], b, self.table, self.geo_params, self.rad_params))
                else:
                    with torch.no_grad():
                        self._sync_if_changed()
                        r, d, m = self.pipe.forward(rays_o[s:e], rays_d[s:e], b)
                        outs.append((r.clone(), d.clone(), m.clone(), self.pipe.counts[:e - s].clone(), self.pipe.t_last[:e - s].clone()))
        finally:
            self.pipe.training = was_training
        rgb, depth, mask, counts, t_last = (torch.cat([o[k] for o in outs], 0) if len(outs) > 1 else outs[0][k] for k in range(5))
        self.last_trans = t_last             # trans_shift[:, -1] of the reference: what FullModel blends a background model with
        # rays without samples: composite already returned T=1 * bkg for the colour; depth takes the far default
        depth = torch.where(counts > 0, depth, torch.full_like(depth, self.depth_far))
        suffix = '' if inference_only else '_coarse'
        return {'rgb' + suffix: rgb, 'depth' + suffix: depth, 'mask' + suffix: mask}

    def train_step(self, rays_o, rays_d, target, bkg=None, next_rays=None):
        """fused alternative to forward + loss + optimiser: one step of the NGP recipe inside the library"""
        self._sync_if_changed()
        loss = self.pipe.train_step(rays_o, rays_d, target, bkg, next_rays=next_rays)
        self._remember_versions()
        return loss
