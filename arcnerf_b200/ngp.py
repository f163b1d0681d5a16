"""The NGP foreground hot path end to end on one GPU, built only from libarcb200 kernels:

    rays -> slab test -> occupancy march (packed samples) -> hash-grid encode -> geometry head (32-64-16)
         -> TruncExp density | [feature, SH(view)] -> radiance head (32-64-64-3, sigmoid) -> compositing
         -> Huber loss -> compositing backward -> radiance/geometry heads backward -> grid scatter -> Adam

This is the configuration of configs/models/nerf_ngp.yaml + configs/expr/NeRF/lego/nerf_lego_nerf_ngp.yaml driven
through the same stages the reference runs (FgModel.forward fg_model.py:153-229 -> NeRF._forward
nerf_model.py:52-91 -> get_sigma_radiance_by_mask_pts fg_model.py:264-318 -> _forward_pts_dir
base_3d_model.py:233-254 -> ray_marching), with two structural differences that do not change results:
  * samples stay PACKED (ray-major list + offsets) instead of being re-expanded to a padded (rays, max_pts) matrix
    and re-compacted with boolean masks; the compositor reproduces the padded-matrix arithmetic;
  * the sample count never leaves the device: every kernel takes it from `meta[0]` (n_dev), so a step is a fixed
    sequence of launches with no host synchronisation and can be replayed as a CUDA graph.
All buffers are allocated once (sized by `capacity` samples and `max_rays` rays).
"""
import math

import numpy as np
import torch

from . import _lib
from . import functional as F

NGP_DEFAULTS = dict(
    n_levels=16, n_features=2, log2_hashmap_size=19, base_res=16, max_res=2048,    # HashGridEmbedder
    geo_width=64, geo_hidden=1, w_feat=16,                                        # FusedMLPGeoNet
    rad_width=64, rad_hidden=2, sh_degree=4,                                      # FusedMLPRadianceNet + SHEmbedder
    n_grid=128, side=2.0, n_sample=1024, near_distance=0.2,                        # VolumeBound / rays
    lr=1e-1, beta1=0.9, beta2=0.999, eps=1e-15, weight_decay=1e-6,                 # optim (Adam)
    huber_delta=0.1, loss_weight=3000.0, trunc_exp_clip=15.0,
)


def per_level_scale(base_res, max_res, n_levels):
    """exp(log(max_res/base_res)/(L-1)) in float32 (hashgrid_encoder.py:81)"""
    return float(torch.exp(torch.log(torch.tensor(max_res / base_res)) / (float(n_levels) - 1)))


def init_params(cfg, seed=1337, grid_scale=1e-4, device='cpu'):
    """Random-init parameters in the reference's layouts: hash table U(-1e-4, 1e-4) (tcnn grid.h:1353-1356), MLP
    matrices xavier-uniform (tcnn fully_fused_mlp.cu:859-885).  Seeded torch CPU generator -> identical bytes for
    the CUDA path and the oracle.  (The reference draws from pcg32(1337); the distributions, not the stream, are
    matched here.)"""
    g = torch.Generator().manual_seed(seed)
    geom = F.GridGeometry(cfg['n_levels'], cfg['n_features'], cfg['log2_hashmap_size'], cfg['base_res'],
                          per_level_scale(cfg['base_res'], cfg['max_res'], cfg['n_levels']))
    table = (torch.rand(geom.n_params, generator=g) * 2 - 1) * grid_scale

    def mlp(in_dim, width, hidden, out_dim):
        in_pad, out_pad = (in_dim + 15) // 16 * 16, (out_dim + 15) // 16 * 16
        shapes = [(width, in_pad)] + [(width, width)] * (hidden - 1) + [(out_pad, width)]
        mats = []
        for o, i in shapes:
            s = math.sqrt(6.0 / (i + o))
            mats.append(((torch.rand(o * i, generator=g) * 2 - 1) * s))
        return torch.cat(mats)

    enc_dim = cfg['n_levels'] * cfg['n_features']
    sh_dim = cfg['sh_degree'] ** 2
    geo = mlp(enc_dim, cfg['geo_width'], cfg['geo_hidden'], cfg['w_feat'])
    rad = mlp(cfg['w_feat'] + sh_dim, cfg['rad_width'], cfg['rad_hidden'], 3)
    return dict(table=table.to(device), geo=geo.to(device), rad=rad.to(device))


class StageTimer:
    """CUDA-event timing of named pipeline stages on the current stream (used by bench.py's roofline pass)."""

    def __init__(self):
        self.records = {}

    def run(self, name, fn):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        out = fn()
        b.record()
        self.records.setdefault(name, []).append((a, b))
        return out

    def summary(self):
        """name -> (mean ms, launches recorded); call after a synchronize"""
        return {k: (sum(a.elapsed_time(b) for a, b in v) / len(v), len(v)) for k, v in self.records.items()}


class NGPPipeline:
    def __init__(self, max_rays=4096, capacity=1 << 19, device='cuda:0', seed=1337, grid_scale=1e-4, params=None,
                 world_size=1, **overrides):
        cfg = dict(NGP_DEFAULTS)
        cfg.update(overrides)
        self.cfg = cfg
        self.dev = torch.device(device)
        self.max_rays, self.capacity = int(max_rays), int(capacity)
        self.world_size = int(world_size)
        _lib.load()
        dev = self.dev
        self.geom = F.GridGeometry(cfg['n_levels'], cfg['n_features'], cfg['log2_hashmap_size'], cfg['base_res'],
                                   per_level_scale(cfg['base_res'], cfg['max_res'], cfg['n_levels']))
        self.enc_dim = self.geom.out_dim
        self.sh_dim = cfg['sh_degree'] ** 2
        self.rad_in_dim = cfg['w_feat'] + self.sh_dim
        self.geo_spec = F.MlpSpec(self.enc_dim, cfg['geo_width'], cfg['geo_hidden'], cfg['w_feat'], 'relu', 'none')
        self.rad_spec = F.MlpSpec(self.rad_in_dim, cfg['rad_width'], cfg['rad_hidden'], 3, 'relu', 'sigmoid')
        if params is None:
            params = init_params(cfg, seed, grid_scale)
        # parameters: fp32 masters (the reference's flat `params` tensors), fp16 table copy, packed MLP operands
        self.table = params['table'].to(dev, torch.float32).contiguous()
        self.geo_p = params['geo'].to(dev, torch.float32).contiguous()
        self.rad_p = params['rad'].to(dev, torch.float32).contiguous()
        assert self.table.numel() == self.geom.n_params
        assert self.geo_p.numel() == self.geo_spec.n_params and self.rad_p.numel() == self.rad_spec.n_params
        self.table_h = F.cast_f16(self.table)
        self.geo_blob = F.mlp_pack(self.geo_p, self.geo_spec)
        self.rad_blob = F.mlp_pack(self.rad_p, self.rad_spec)
        z = lambda t: torch.zeros_like(t)
        self.g_table, self.m_table, self.v_table = z(self.table), z(self.table), z(self.table)
        self.g_geo, self.m_geo, self.v_geo = z(self.geo_p), z(self.geo_p), z(self.geo_p)
        self.g_rad, self.m_rad, self.v_rad = z(self.rad_p), z(self.rad_p), z(self.rad_p)
        self.step_count = 0
        # occupancy + box
        n = cfg['n_grid']
        half = cfg['side'] / 2.0
        self.aabb = torch.tensor([[-half] * 3, [half] * 3], dtype=torch.float32, device=dev)   # (2,3) min; max
        self.volume = torch.ones((n, n, n), dtype=torch.bool, device=dev)
        self.dt = float(np.float32(cfg['side'] * math.sqrt(3.0)) / np.float32(cfg['n_sample']))   # diag / n_sample
        self.rng = _lib.Rng()
        self.timer = None
        self._alloc()

    def _run(self, name, fn):
        return fn() if self.timer is None else self.timer.run(name, fn)

    # ------------------------------------------------------------------------------------------ buffers
    def _alloc(self):
        dev, R, C = self.dev, self.max_rays, self.capacity
        f32, i32, f16 = torch.float32, torch.int32, torch.float16
        e = lambda shape, dt: torch.empty(shape, dtype=dt, device=dev)
        self.near, self.far, self.hit = e((R, 1), f32), e((R, 1), f32), e((R, 1), torch.uint8)
        self.counts, self.offsets, self.meta = e((R,), i32), e((R + 1,), i32), torch.zeros(2, dtype=i32, device=dev)
        nbytes = _lib.load().arcb200_march_packed_scratch_bytes(R, self.cfg['n_sample'])
        self.scratch = e((nbytes,), torch.uint8)
        self.t, self.unit, self.ray_id = e((C,), f32), e((C, 3), f32), e((C,), i32)
        self.enc = e((C, self.enc_dim), f16)
        self.rad_in = torch.zeros((C, self.rad_in_dim), dtype=f16, device=dev)
        self.sigma, self.rgb_s = e((C,), f32), e((C, 3), f32)
        self.rgb, self.depth, self.acc = e((R, 3), f32), e((R,), f32), e((R,), f32)
        self.g_rgb = e((R, 3), f32)
        self.d_sigma, self.d_rgb_s = torch.zeros((C,), dtype=f32, device=dev), torch.zeros((C, 3), dtype=f32, device=dev)
        self.d_rad_in = e((C, self.rad_in_dim), f32)
        self.d_geo_out = e((C, self.cfg['w_feat']), f32)
        self.d_enc = e((C, self.enc_dim), f32)
        self.loss = torch.zeros(1, dtype=f32, device=dev)
        self.n_dev = self.meta[0:1]
        self.pmax = self.meta[1:2]

    def set_occupancy(self, volume_bool):
        """(n,n,n) bool occupancy (numpy or tensor), the VolumeBound bitfield."""
        v = torch.as_tensor(volume_bool).to(self.dev, torch.bool).contiguous()
        assert v.shape == self.volume.shape
        self.volume.copy_(v)

    # ------------------------------------------------------------------------------------------ stages
    def sample(self, rays_o, rays_d):
        """slab test + packed occupancy march; everything stays on the device"""
        cfg = self.cfg
        R = rays_o.shape[0]
        assert R <= self.max_rays
        st = _lib.current_stream()
        self._run('aabb', lambda: _lib.call(
            'arcb200_aabb_intersection', rays_o.data_ptr(), rays_d.data_ptr(), self.aabb.data_ptr(),
            self.near.data_ptr(), self.far.data_ptr(), None, self.hit.data_ptr(), R, 1, st))
        self._run('march_packed', lambda: _lib.call(
            'arcb200_march_packed', 0, rays_o.data_ptr(), rays_d.data_ptr(), self.near.data_ptr(),
            self.far.data_ptr(), cfg['n_sample'], self.dt, self.aabb.data_ptr(), cfg['n_grid'],
            self.volume.data_ptr(), cfg['near_distance'], R, self.rng.state, self.rng.inc,
            self.scratch.data_ptr(), self.counts.data_ptr(), self.offsets.data_ptr(), self.meta.data_ptr(),
            self.capacity, self.t.data_ptr(), None, self.ray_id.data_ptr(), self.unit.data_ptr(), st))
        self.rng.advance()

    def field_forward(self, rays_d):
        """encode -> geometry head -> density | [feat, SH] -> radiance head, on the packed samples"""
        C, nd = self.capacity, self.n_dev
        self._run('grid_fwd', lambda: F.grid_forward(self.unit, self.table_h, self.geom, n_dev=nd, out=self.enc))
        # geometry head writes its W_feat outputs straight into the radiance input row; SH fills the rest
        self._run('mlp_geo_fwd', lambda: F.mlp_forward(self.enc, self.geo_blob, self.geo_spec, out=self.rad_in, out_col=0, n_dev=nd))
        self._run('sh_fwd', lambda: F.sh_forward(rays_d, self.cfg['sh_degree'], index=self.ray_id, remap=True, out=self.rad_in,
                                                 out_col=self.cfg['w_feat'], n_dev=nd))
        self._run('trunc_exp_fwd', lambda: _lib.call(
            'arcb200_trunc_exp_fwd', self.rad_in.data_ptr(), 1, self.rad_in.stride(0), self.sigma.data_ptr(), C,
            nd.data_ptr(), _lib.current_stream()))
        self._run('mlp_rad_fwd', lambda: F.mlp_forward(self.rad_in, self.rad_blob, self.rad_spec, out=self.rgb_s, n_dev=nd))

    def composite(self, R, bkg):
        bkg_per_ray = int(bkg is not None and bkg.shape[0] == R and R > 1)
        self._run('composite_fwd', lambda: _lib.call(
            'arcb200_composite_fwd', self.offsets.data_ptr(), 0, self.capacity, self.pmax.data_ptr(),
            self.sigma.data_ptr(), 0, self.rgb_s.data_ptr(), self.t.data_ptr(), None, 0, 0, _lib.ptr(bkg),
            bkg_per_ray, self.rgb.data_ptr(), self.depth.data_ptr(), self.acc.data_ptr(), None, None, None, R,
            _lib.current_stream()))

    def forward(self, rays_o, rays_d, bkg=None):
        """-> rgb (R,3), depth (R,), mask (R,) views of the pipeline's output buffers"""
        R = rays_o.shape[0]
        self.sample(rays_o, rays_d)
        self.field_forward(rays_d)
        self.composite(R, bkg)
        return self.rgb[:R], self.depth[:R], self.acc[:R]

    def backward(self, rays_d, R, bkg):
        cfg, C, nd, st = self.cfg, self.capacity, self.n_dev, _lib.current_stream()
        bkg_per_ray = int(bkg is not None and bkg.shape[0] == R and R > 1)
        self._run('composite_bwd', lambda: _lib.call(
            'arcb200_composite_bwd', self.offsets.data_ptr(), 0, C, self.pmax.data_ptr(), self.sigma.data_ptr(), 0,
            self.rgb_s.data_ptr(), self.t.data_ptr(), None, 0, 0, _lib.ptr(bkg), bkg_per_ray,
            self.g_rgb.data_ptr(), None, None, None, self.d_sigma.data_ptr(), self.d_rgb_s.data_ptr(),
            cfg['n_sample'], R, st))
        self._run('mlp_rad_bwd', lambda: F.mlp_backward(self.rad_in, self.rad_blob, self.rad_spec, self.d_rgb_s,
                                                        dW=self.g_rad, dx=self.d_rad_in, n_dev=nd))
        self._run('geo_head_grad', lambda: _lib.call(
            'arcb200_geo_head_grad', self.rad_in.data_ptr(), 1, self.rad_in.stride(0), self.d_sigma.data_ptr(),
            cfg['trunc_exp_clip'], self.d_rad_in.data_ptr(), 0, self.d_rad_in.stride(0), cfg['w_feat'],
            self.d_geo_out.data_ptr(), C, nd.data_ptr(), st))
        self._run('mlp_geo_bwd', lambda: F.mlp_backward(self.enc, self.geo_blob, self.geo_spec, self.d_geo_out,
                                                        dW=self.g_geo, dx=self.d_enc, n_dev=nd))
        self._run('grid_bwd', lambda: F.grid_backward(self.unit, self.d_enc, self.geom, grad_table=self.g_table, n_dev=nd))

    def optimizer_step(self):
        cfg = self.cfg
        self.step_count += 1
        gs = 1.0 / self.world_size
        kw = dict(beta1=cfg['beta1'], beta2=cfg['beta2'], eps=cfg['eps'], weight_decay=cfg['weight_decay'], grad_scale=gs)
        self._run('adam_table', lambda: F.adam_step(self.table, self.m_table, self.v_table, self.g_table, self.step_count,
                                                    cfg['lr'], p_half=self.table_h, **kw))

        def mlp_update():
            F.adam_step(self.geo_p, self.m_geo, self.v_geo, self.g_geo, self.step_count, cfg['lr'], **kw)
            F.adam_step(self.rad_p, self.m_rad, self.v_rad, self.g_rad, self.step_count, cfg['lr'], **kw)
            F.mlp_pack(self.geo_p, self.geo_spec, self.geo_blob)
            F.mlp_pack(self.rad_p, self.rad_spec, self.rad_blob)
        self._run('adam_mlp+pack', mlp_update)

    def all_reduce_grads(self):
        """Data-parallel exchange: sum of the hash-table and MLP gradients over ranks (NCCL over NVLink); the
        reference gets the same from DDP's bucketed all-reduce (basic_trainer.py:192-202)."""
        if self.world_size > 1:
            from .dist import all_reduce_sum_
            all_reduce_sum_([self.g_table, self.g_geo, self.g_rad])

    def train_step(self, rays_o, rays_d, target, bkg=None, optimize=True):
        """One optimisation step on a ray batch; returns the (device) loss tensor.  No host synchronisation."""
        R = rays_o.shape[0]
        self.forward(rays_o, rays_d, bkg)
        self.loss.zero_()
        self._run('huber_loss', lambda: _lib.call(
            'arcb200_huber_loss', self.rgb.data_ptr(), target.data_ptr(), self.cfg['huber_delta'],
            self.cfg['loss_weight'], self.g_rgb.data_ptr(), self.loss.data_ptr(), R, _lib.current_stream()))
        self.backward(rays_d, R, bkg)
        if optimize:
            self._run('all_reduce', self.all_reduce_grads)
            self.optimizer_step()
        return self.loss

    @torch.no_grad()
    def render(self, rays_o, rays_d, bkg=None, chunk=None):
        """Full-frame forward in chunks of `max_rays` rays (the reference chunks by chunk_rays=32768,
        full_model.py:415-457); returns rgb (N,3), depth (N,), mask (N,)."""
        N = rays_o.shape[0]
        chunk = chunk or self.max_rays
        rgb = torch.empty((N, 3), dtype=torch.float32, device=self.dev)
        depth = torch.empty((N,), dtype=torch.float32, device=self.dev)
        acc = torch.empty((N,), dtype=torch.float32, device=self.dev)
        for s in range(0, N, chunk):
            e = min(N, s + chunk)
            b = None if bkg is None else (bkg if bkg.shape[0] == 1 else bkg[s:e])
            r, d, a = self.forward(rays_o[s:e], rays_d[s:e], b)
            rgb[s:e].copy_(r); depth[s:e].copy_(d); acc[s:e].copy_(a)
        return rgb, depth, acc

    # number of library kernels one train_step enqueues (for bench.py's gpu_launches)
    LAUNCHES_FORWARD = 1 + 4 + 1 + 1 + 1 + 1 + 1 + 1      # aabb, march(4), grid, geo mlp, sh, exp, rad mlp, composite
    LAUNCHES_TRAIN = LAUNCHES_FORWARD + 1 + 1 + 1 + 1 + 1 + 1 + 3 + 2   # huber, comp bwd, rad bwd, glue, geo bwd, grid bwd, adam x3, pack x2
