"""CPU oracle of the whole NGP foreground step, composed the way the reference composes it (padded (rays, max_pts)
matrices, boolean-mask compaction).  TEST INFRASTRUCTURE ONLY: used by tests/, by
__graft_entry__.smoke() and, timed, by bench.py's cpu_baseline / --impl reference legs.

Stages and the reference lines they restate:
  slab test + occupancy march      oracle/sampler_oracle.c (arcnerf/ops CUDA kernels)
  reduce_empty_mask / padding      arcnerf/models/fg_model.py:251-262, 264-318
  hash-grid encode, SH, fused MLPs tiny-cuda-nn arithmetic (oracle/fields.py); a torch-CPU (multi-threaded) copy of
                                   the same arithmetic lives here for the timed baseline and is cross-checked
                                   against the numpy one in tests/test_oracle_ngp.py
  TruncExp density                 arcnerf/ops/trunc_exp.py:7-37
  ray_marching                     arcnerf/render/ray_helper.py:476-620 (oracle/render.py)
  Huber image loss x3000           arcnerf/loss/img_loss.py:48-98
  Adam                             torch.optim.Adam (common/trainer/optimizer.py)
"""
import numpy as np
import torch

from . import fields, render
from .sampler import Sampler

PY = 2654435761
PZ = 805459861

# The reference's NGP configuration (configs/models/nerf_ngp.yaml + configs/expr/NeRF/lego/nerf_lego_nerf_ngp.yaml); the same
# values as arcnerf_b200.ngp.NGP_DEFAULTS (tests/test_oracle_ngp.py checks that), kept here so that the timed CPU arm of
# bench.py never imports the product package.
NGP_CONFIG = dict(
    n_levels=16, n_features=2, log2_hashmap_size=19, base_res=16, max_res=2048,
    geo_width=64, geo_hidden=1, w_feat=16, rad_width=64, rad_hidden=2, sh_degree=4,
    n_grid=128, side=2.0, n_sample=1024, near_distance=0.2,
    lr=1e-1, beta1=0.9, beta2=0.999, eps=1e-15, weight_decay=1e-6,
    huber_delta=0.1, loss_weight=3000.0, trunc_exp_clip=15.0,
)


def init_params(cfg, seed=1337, grid_scale=1e-4):
    """Random-init parameters in the reference's layouts, byte-identical to arcnerf_b200.ngp.init_params (same torch CPU
    generator, same draw order) but with the level geometry taken from oracle.fields.grid_geometry: hash table
    U(-grid_scale, grid_scale) (tcnn grid.h:1353-1356), MLP matrices xavier-uniform (fully_fused_mlp.cu:859-885)."""
    import math
    g = torch.Generator().manual_seed(seed)
    pls = float(torch.exp(torch.log(torch.tensor(cfg['max_res'] / cfg['base_res'])) / (float(cfg['n_levels']) - 1)))
    geom = fields.grid_geometry(cfg['n_levels'], cfg['log2_hashmap_size'], cfg['base_res'], pls)
    n_table = int(geom[0][-1]) * cfg['n_features']
    table = (torch.rand(n_table, generator=g) * 2 - 1) * grid_scale

    def mlp(in_dim, width, hidden, out_dim):
        in_pad, out_pad = (in_dim + 15) // 16 * 16, (out_dim + 15) // 16 * 16
        shapes = [(width, in_pad)] + [(width, width)] * (hidden - 1) + [(out_pad, width)]
        return torch.cat([(torch.rand(o * i, generator=g) * 2 - 1) * math.sqrt(6.0 / (i + o)) for o, i in shapes])

    geo = mlp(cfg['n_levels'] * cfg['n_features'], cfg['geo_width'], cfg['geo_hidden'], cfg['w_feat'])
    rad = mlp(cfg['w_feat'] + cfg['sh_degree'] ** 2, cfg['rad_width'], cfg['rad_hidden'], 3)
    return dict(table=table, geo=geo, rad=rad)


class TorchGrid:
    """torch-CPU restatement of the hash-grid encode / table gradient (same arithmetic as fields.grid_encode with
    accumulate='fp32'), vectorised over samples so that it uses all host threads."""

    def __init__(self, geom):
        self.offsets, self.scales, self.res, self.hashed = [np.asarray(a) for a in geom[:4]]
        self.L = len(self.scales)

    def _corners(self, x, l):
        scale = float(self.scales[l])
        pos = (x.double() * scale + 0.5).float()          # fma emulation: one rounding
        fl = torch.floor(pos)
        g = fl.to(torch.int64) & 0xFFFFFFFF
        f = pos - fl
        size, r, off = int(self.offsets[l + 1] - self.offsets[l]), int(self.res[l]), int(self.offsets[l])
        out = []
        for k in range(8):
            c = (k & 1, (k >> 1) & 1, (k >> 2) & 1)
            gx, gy, gz = ((g[:, d] + c[d]) & 0xFFFFFFFF for d in range(3))
            if self.hashed[l]:
                idx = (gx ^ ((gy * PY) & 0xFFFFFFFF) ^ ((gz * PZ) & 0xFFFFFFFF)) & 0xFFFFFFFF
            else:
                idx = (gx + ((gy * r) & 0xFFFFFFFF) + ((gz * ((r * r) & 0xFFFFFFFF)) & 0xFFFFFFFF)) & 0xFFFFFFFF
            idx = idx % size + off
            w = ((f[:, 0] if c[0] else 1 - f[:, 0]) * (f[:, 1] if c[1] else 1 - f[:, 1])) * (f[:, 2] if c[2] else 1 - f[:, 2])
            out.append((idx, w))
        return out

    def encode(self, x, table_h, F=2):
        """x (N,3) float32; table_h (entries, F) float32 holding fp16 values -> (N, L*F) float32 (fp16 values)"""
        outs = []
        for l in range(self.L):
            acc = torch.zeros((x.shape[0], F), dtype=torch.float32)
            for idx, w in self._corners(x, l):
                acc = acc + w[:, None] * table_h[idx]
            outs.append(acc.half().float())
        return torch.cat(outs, 1)

    def backward(self, x, d_enc, n_entries, F=2):
        grad = torch.zeros((n_entries, F), dtype=torch.float32)
        for l in range(self.L):
            g = d_enc[:, l * F:(l + 1) * F]
            for idx, w in self._corners(x, l):
                grad.index_add_(0, idx, g * w[:, None])
        return grad


def _mlp_mats(params, dims):
    return [torch.from_numpy(m.astype(np.float16).astype(np.float32)) for m in fields.mlp_split(np.asarray(params), *dims)]


def mlp_forward_t(x, mats, hidden_act=torch.relu, out_act=None, out_dim=None, in_pad=None):
    """fp16 weights/activations, fp32 accumulate, torch CPU.  Returns (out (N,out_dim), list of layer inputs)."""
    n, d = x.shape
    h = torch.ones((n, in_pad), dtype=torch.float32)
    h[:, :d] = x.half().float()
    acts = [h]
    for l, w in enumerate(mats):
        z = h @ w.t()
        if l < len(mats) - 1:
            h = hidden_act(z).half().float()
            acts.append(h)
        else:
            y = out_act(z) if out_act is not None else z
    return y[:, :out_dim], acts, y


def mlp_backward_t(acts, y, mats, dy, out_sigmoid, in_dim):
    g = torch.zeros_like(y)
    g[:, :dy.shape[1]] = dy * (y[:, :dy.shape[1]] * (1 - y[:, :dy.shape[1]]) if out_sigmoid else 1.0)
    dws = [None] * len(mats)
    for l in range(len(mats) - 1, -1, -1):
        dws[l] = g.t() @ acts[l]
        g = g @ mats[l]
        if l > 0:
            g = g * (acts[l] > 0).float()
    return g[:, :in_dim], torch.cat([d.reshape(-1) for d in dws])


class NGPOracle:
    def __init__(self, params, cfg, threads=None):
        self.cfg = dict(cfg)
        self.smp = Sampler('oracle')
        if threads:
            self.smp.lib.orc_set_threads(int(threads))
        pls = float(torch.exp(torch.log(torch.tensor(cfg['max_res'] / cfg['base_res'])) / (float(cfg['n_levels']) - 1)))
        self.geom = fields.grid_geometry(cfg['n_levels'], cfg['log2_hashmap_size'], cfg['base_res'], pls)
        self.tg = TorchGrid(self.geom)
        self.F = cfg['n_features']
        self.enc_dim = cfg['n_levels'] * self.F
        self.sh_dim = cfg['sh_degree'] ** 2
        self.geo_dims = (self.enc_dim, cfg['geo_width'], cfg['geo_hidden'], cfg['w_feat'])
        self.rad_dims = (cfg['w_feat'] + self.sh_dim, cfg['rad_width'], cfg['rad_hidden'], 3)
        self.table = np.asarray(params['table'], np.float32).copy()
        self.geo_p = np.asarray(params['geo'], np.float32).copy()
        self.rad_p = np.asarray(params['rad'], np.float32).copy()
        half = cfg['side'] / 2.0
        self.aabb = np.array([[-half] * 3, [half] * 3], np.float32)
        self.volume = np.ones((cfg['n_grid'],) * 3, bool)
        self.dt = np.float32(np.float32(cfg['side'] * np.sqrt(3.0)) / np.float32(cfg['n_sample']))
        self.rng = self.smp.rng()
        self.adam = None
        self.step_count = 0

    def set_occupancy(self, v):
        self.volume = np.asarray(v, bool)

    def sample(self, o, d):
        cfg = self.cfg
        near, far, _, hit = self.smp.aabb_intersection(o, d, self.aabb[None])
        z, m = self.smp.sparse_volume_sampling(o, d, near, far, cfg['n_sample'], self.dt, self.aabb, cfg['n_grid'],
                                               self.volume, cfg['near_distance'], self.rng)
        self.rng.advance()
        # reduce_empty_mask (fg_model.py:251-262)
        p = max(2, int(m.sum(1).max()))
        return z[:, :p], m[:, :p]

    def field(self, pts_unit, dirs):
        """(N,3) unit-cube positions, (N,3) view dirs -> sigma (N,), rgb (N,3), plus what the backward needs"""
        cfg = self.cfg
        table_h = torch.from_numpy(self.table.astype(np.float16).astype(np.float32).reshape(-1, self.F))
        enc = self.tg.encode(torch.from_numpy(pts_unit), table_h, self.F)
        geo_mats = _mlp_mats(self.geo_p, self.geo_dims)
        rad_mats = _mlp_mats(self.rad_p, self.rad_dims)
        geo_out, geo_acts, geo_y = mlp_forward_t(enc, geo_mats, out_dim=cfg['w_feat'], in_pad=(self.enc_dim + 15) // 16 * 16)
        geo_out = geo_out.half().float()                        # the head's output is stored in fp16
        sh = torch.from_numpy(fields.sh_encode(dirs, cfg['sh_degree'], remap=True)).half().float()
        rad_in = torch.cat([geo_out, sh], 1)
        rgb, rad_acts, rad_y = mlp_forward_t(rad_in, rad_mats, out_act=torch.sigmoid, out_dim=3,
                                             in_pad=(self.rad_dims[0] + 15) // 16 * 16)
        sigma = torch.exp(geo_out[:, 0])
        ctx = dict(enc=enc, geo_mats=geo_mats, rad_mats=rad_mats, geo_acts=geo_acts, geo_y=geo_y, rad_acts=rad_acts,
                   rad_y=rad_y, geo_out=geo_out, pts_unit=pts_unit)
        return sigma, rgb, ctx

    def forward(self, o, d, bkg=None, keep=False):
        """-> dict(rgb, depth, mask, counts, [ctx])  following FgModel.forward / NeRF._forward on padded matrices"""
        z, m = self.sample(o, d)
        counts = m.sum(1)
        r, p = z.shape
        pts = o[:, None, :] + d[:, None, :] * z[:, :, None]     # get_ray_points_by_zvals
        pts_v = pts[m].astype(np.float32)
        dirs_v = np.repeat(d[:, None, :], p, 1)[m]
        unit = ((pts_v - self.aabb[0]) / (self.aabb[1] - self.aabb[0])).astype(np.float32)
        sigma_v, rgb_v, ctx = self.field(unit, dirs_v)
        sigma_v = sigma_v.clone().requires_grad_(keep)
        rgb_v = rgb_v.clone().requires_grad_(keep)
        # scatter back to dense with last-sample padding (fg_model.py:311-316)
        mt = torch.from_numpy(m)
        sigma = torch.zeros((r, p)); rad = torch.zeros((r, p, 3))
        if mt.any():
            last = torch.from_numpy(np.cumsum(counts) - 1).clamp(min=0)
            sigma = torch.ones((r, p)) * sigma_v[last].unsqueeze(1)
            rad = torch.ones((r, p, 3)) * rgb_v[last].unsqueeze(1)
            sigma = sigma.masked_scatter(mt, sigma_v)
            rad = rad.masked_scatter(mt.unsqueeze(-1).expand(-1, -1, 3), rgb_v)
        valid = torch.from_numpy(counts > 0)
        out = render.ray_marching(sigma, rad, torch.from_numpy(z), bkg_color=None if bkg is None else torch.from_numpy(bkg))
        rgb = out['rgb']
        if bkg is not None:      # update_values_for_invalid_rays: background colour for rays without samples
            b = torch.from_numpy(bkg).expand(r, 3)
            rgb = torch.where(valid[:, None], rgb, b)
        res = dict(rgb=rgb, depth=out['depth'] * valid, mask=out['mask'] * valid, counts=counts.astype(np.int32),
                   z=z, m=m)
        if keep:
            res.update(ctx=ctx, sigma_v=sigma_v, rgb_v=rgb_v)
        return res

    def loss_and_grads(self, o, d, target, bkg=None):
        """-> loss (float), dict of parameter gradients in the flat layouts"""
        cfg = self.cfg
        res = self.forward(o, d, bkg, keep=True)
        loss = render.huber_loss(res['rgb'], torch.from_numpy(target), cfg['huber_delta'], cfg['loss_weight'])
        if res['sigma_v'].numel() == 0:
            return float(loss.detach()), dict(table=np.zeros_like(self.table), geo=np.zeros_like(self.geo_p), rad=np.zeros_like(self.rad_p)), res
        loss.backward()
        ctx = res['ctx']
        d_sigma, d_rgb = res['sigma_v'].grad, res['rgb_v'].grad
        d_rad_in, dw_rad = mlp_backward_t(ctx['rad_acts'], ctx['rad_y'], ctx['rad_mats'], d_rgb, True, self.rad_dims[0])
        d_geo = d_rad_in[:, :cfg['w_feat']].clone()
        d_geo[:, 0] += d_sigma * torch.exp(ctx['geo_out'][:, 0].clamp(-cfg['trunc_exp_clip'], cfg['trunc_exp_clip']))
        d_enc, dw_geo = mlp_backward_t(ctx['geo_acts'], ctx['geo_y'], ctx['geo_mats'], d_geo, False, self.enc_dim)
        g_table = self.tg.backward(torch.from_numpy(ctx['pts_unit']), d_enc, self.table.size // self.F, self.F)
        return float(loss.detach()), dict(table=g_table.reshape(-1).numpy(), geo=dw_geo.numpy(), rad=dw_rad.numpy()), res

    def train_step(self, o, d, target, bkg=None):
        cfg = self.cfg
        loss, grads, _ = self.loss_and_grads(o, d, target, bkg)
        if self.adam is None:
            self.tensors = {k: torch.from_numpy(getattr(self, n)).requires_grad_(True)
                            for k, n in (('table', 'table'), ('geo', 'geo_p'), ('rad', 'rad_p'))}
            self.adam = torch.optim.Adam(list(self.tensors.values()), lr=cfg['lr'], betas=(cfg['beta1'], cfg['beta2']),
                                         eps=cfg['eps'], weight_decay=cfg['weight_decay'])
        for k, t in self.tensors.items():
            t.grad = torch.from_numpy(np.ascontiguousarray(grads[k], np.float32))
        self.adam.step()
        self.step_count += 1
        return loss
