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
import os

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


class SampleSet:
    """Everything the sampler produces for one ray batch (packed layout).  The pipeline owns two of them
    so that the march of the next batch can run on a side stream while the current batch is in the field kernels."""

    def __init__(self, max_rays, capacity, n_sample, dev):
        f32, i32 = torch.float32, torch.int32
        e = lambda shape, dt: torch.empty(shape, dtype=dt, device=dev)
        R, C = max_rays, capacity
        self.near, self.far, self.hit = e((R, 1), f32), e((R, 1), f32), e((R, 1), torch.uint8)
        self.counts, self.offsets = e((R,), i32), e((R + 1,), i32)
        self.meta = torch.zeros(2, dtype=i32, device=dev)
        self.scratch = e((_lib.load().arcb200_march_packed_scratch_bytes(R, n_sample),), torch.uint8)
        self.t, self.unit, self.ray_id = e((C,), f32), e((C, 3), f32), e((C,), i32)
        self.n_dev, self.pmax = self.meta[0:1], self.meta[1:2]
        self.ready = None        # event recorded after the sampler kernels (side stream) or None
        self.key = None          # (rays_o.data_ptr(), rays_d.data_ptr(), n_rays) the set was filled for


class NGPPipeline:
    def __init__(self, max_rays=4096, capacity=1 << 19, device='cuda:0', seed=1337, grid_scale=1e-4, params=None,
                 world_size=1, rank=0, exchange='auto', **overrides):
        """world_size > 1: data-parallel over the ranks of torch.distributed's default group.  exchange = 'peer' (the
        gradient exchange fused into the Adam kernel over NVLink peer memory, csrc/peer.cu; ranks on one node),
        'peer16' (the same with the hash-table gradient travelling as a loss-scaled fp16 copy - the precision tcnn's
        gradients have when the reference hands them to DDP - and summed in fp32), 'nccl' (one all-reduce of the flat
        gradient buffer, then the single-GPU Adam), or 'auto' (= 'peer16' unless ARCB200_EXCHANGE says otherwise)."""
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
        nt, ng, nr = self.table.numel(), self.geo_p.numel(), self.rad_p.numel()
        self.rank = int(rank)
        if exchange == 'auto':
            exchange = os.environ.get('ARCB200_EXCHANGE', 'peer16')
        self.exchange = exchange if self.world_size > 1 else 'none'
        self.arena = None
        if self.exchange in ('peer', 'peer16', 'mc16'):
            import torch.distributed as dist
            from .peer import PeerArena, SymmArena, slice_bounds

            def share(handle):
                out = [None] * self.world_size
                dist.all_gather_object(out, handle)
                return out

            def all_ok(flag):
                ok = torch.tensor([1 if flag else 0], dtype=torch.int32, device=dev if dist.get_backend() != 'gloo' else 'cpu')
                dist.all_reduce(ok, op=dist.ReduceOp.MIN)
                return int(ok.item()) == 1

            def note(msg):
                if self.rank == 0:
                    import sys
                    sys.stderr.write('arcnerf_b200: %s\n' % msg)
            self.rank = dist.get_rank()
            regions = {'grad': (nt + ng + nr) * 4, 'table_h': nt * 2}
            if self.exchange in ('peer16', 'mc16'):
                regions['grad16'] = nt * 2
            if self.exchange == 'mc16':
                # in-switch reduce / broadcast: needs torch's symmetric memory with a multicast mapping on every rank (NVSwitch)
                err = None
                try:
                    assert nt % 8 == 0, 'table size not a multiple of 8'
                    self.arena = SymmArena(regions, dev, self.rank, self.world_size)
                    if not self.arena.mc_base:
                        err = 'no multicast support on this fabric'
                except Exception as e:                          # noqa: BLE001  (any failure -> agree on the fallback)
                    err = e
                if not all_ok(err is None):
                    note('multicast exchange unavailable (%s); using the peer-memory exchange' % (err if err is not None else 'another rank failed'))
                    if self.arena is not None:
                        self.arena.close(sync_ranks=dist.barrier)
                    else:
                        dist.barrier()
                    self.arena, self.exchange = None, 'peer16'
            if self.arena is None:
                # peer mapping needs every rank to see every GPU of the node (a launcher that hides devices per process, or GPUs
                # without P2P, makes cudaIpcOpenMemHandle fail): all ranks then agree to use the NCCL all-reduce instead
                err = None
                try:
                    self.arena = PeerArena(regions, dev, self.rank, self.world_size, share)
                except RuntimeError as e:
                    err = e
                if not all_ok(err is None):
                    note('peer-memory exchange unavailable (%s); using the NCCL all-reduce'
                         % (err if err is not None else 'another rank could not map the arenas'))
                    if self.arena is not None:   # mapped here, but another rank could not: give the mappings and the arena back
                        self.arena.close(sync_ranks=dist.barrier)
                    else:
                        dist.barrier()
                    self.arena, self.exchange = None, 'nccl'
        if self.arena is not None:
            if os.environ.get('ARCB200_PEER_HOST_BARRIER', '0') == '1':
                self.arena.host_barrier = dist.barrier          # ranks sharing one device (tests): no spinning kernels
            self.table_h = self.arena.local('table_h', torch.float16)
            F.cast_f16(self.table, self.table_h)
            self.slice = slice_bounds(nt, self.rank, self.world_size)
            dist.barrier()                                  # every arena is mapped and initialised before anyone steps
        else:
            self.table_h = F.cast_f16(self.table)
        self.geo_blob = F.mlp_pack(self.geo_p, self.geo_spec)
        self.rad_blob = F.mlp_pack(self.rad_p, self.rad_spec)
        z = lambda t: torch.zeros_like(t)
        # one flat gradient buffer (hash table | geometry head | radiance head): a single collective moves it all
        if self.arena is not None:
            self.g_all = self.arena.local('grad', torch.float32)
            self.g_all.zero_()
        else:
            self.g_all = torch.zeros(nt + ng + nr, dtype=torch.float32, device=dev)
        self.g_table, self.g_geo, self.g_rad = self.g_all[:nt], self.g_all[nt:nt + ng], self.g_all[nt + ng:]
        self.m_table, self.v_table = z(self.table), z(self.table)
        self.m_geo, self.v_geo = z(self.geo_p), z(self.geo_p)
        self.m_rad, self.v_rad = z(self.rad_p), z(self.rad_p)
        self.step_count = 0
        self.step_dev = torch.zeros(1, dtype=torch.int32, device=dev)     # same count, device resident (graph replays)
        # occupancy + box
        n = cfg['n_grid']
        half = cfg['side'] / 2.0
        self.aabb = torch.tensor([[-half] * 3, [half] * 3], dtype=torch.float32, device=dev)   # (2,3) min; max
        self.volume = torch.ones((n, n, n), dtype=torch.bool, device=dev)
        # occupancy='bitfield' (overrides): march through the Morton-ordered bit field of BitfieldBound instead of the bool
        # volume of VolumeBound (sparse_volume_sampling_bit, arcnerf/ops/bitfield_func.py:17-61); 256 KB instead of 2 MB
        self.use_bits = cfg.get('occupancy', 'volume') == 'bitfield'
        self.bits = torch.full((n * n * n // 8, ), 255, dtype=torch.uint8, device=dev) if self.use_bits else None
        self.dt = float(np.float32(cfg['side'] * math.sqrt(3.0)) / np.float32(cfg['n_sample']))   # diag / n_sample
        self.rng = _lib.Rng()
        self.timer = None
        # the fused field forward covers exactly the reference's NGP configuration (ARCB200_FUSED_FWD=0: separate launches)
        self.fused_fwd = (os.environ.get('ARCB200_FUSED_FWD', '1') == '1' and cfg['n_levels'] == 16 and cfg['n_features'] == 2
                          and (cfg['geo_width'], cfg['geo_hidden'], cfg['w_feat']) == (64, 1, 16)
                          and (cfg['rad_width'], cfg['rad_hidden'], cfg['sh_degree']) == (64, 2, 4))
        self._alloc()

    def _run(self, name, fn):
        return fn() if self.timer is None else self.timer.run(name, fn)

    # ------------------------------------------------------------------------------------------ buffers
    def _alloc(self):
        dev, R, C = self.dev, self.max_rays, self.capacity
        f32, i32, f16 = torch.float32, torch.int32, torch.float16
        e = lambda shape, dt: torch.empty(shape, dtype=dt, device=dev)
        self.sets = [SampleSet(R, C, self.cfg['n_sample'], dev), SampleSet(R, C, self.cfg['n_sample'], dev)]
        self.cur = 0
        self.forward_serial = 0     # number of sample sets made current so far (stale-activation guard of the autograd wrapper)
        self.side_stream = torch.cuda.Stream(device=dev)
        self.opt_stream = torch.cuda.Stream(device=dev)
        self._opt_fork, self._opt_join = torch.cuda.Event(), torch.cuda.Event()
        self.overlap_heads = os.environ.get('ARCB200_OVERLAP_HEADS', '1') == '1'
        self.enc = e((C, self.enc_dim), f16)
        self.rad_in = torch.zeros((C, self.rad_in_dim), dtype=f16, device=dev)
        self.sigma, self.rgb_s = e((C,), f32), e((C, 3), f32)
        self.rgb, self.depth, self.acc = e((R, 3), f32), e((R,), f32), e((R,), f32)
        self.t_last = e((R,), f32)          # transmittance the background is weighted with (trans_shift[:, -1])
        self.g_rgb = e((R, 3), f32)
        self.d_sigma, self.d_rgb_s = torch.zeros((C,), dtype=f32, device=dev), torch.zeros((C, 3), dtype=f32, device=dev)
        self.d_rad_in = e((C, self.rad_in_dim), f32)
        # hidden activations handed from the forward to the backward kernels (operand-tile layout)
        self.geo_act = F.mlp_act_save_buffer(C, self.geo_spec, dev)
        self.rad_act = F.mlp_act_save_buffer(C, self.rad_spec, dev)
        self.training = True
        self.d_enc = e((C, self.enc_dim), f32)
        self.loss = torch.zeros(1, dtype=f32, device=dev)

    # the sampler outputs of the batch currently in the field kernels
    near = property(lambda self: self.sets[self.cur].near)
    far = property(lambda self: self.sets[self.cur].far)
    hit = property(lambda self: self.sets[self.cur].hit)
    counts = property(lambda self: self.sets[self.cur].counts)
    offsets = property(lambda self: self.sets[self.cur].offsets)
    meta = property(lambda self: self.sets[self.cur].meta)
    t = property(lambda self: self.sets[self.cur].t)
    unit = property(lambda self: self.sets[self.cur].unit)
    ray_id = property(lambda self: self.sets[self.cur].ray_id)
    n_dev = property(lambda self: self.sets[self.cur].n_dev)
    pmax = property(lambda self: self.sets[self.cur].pmax)

    def sync_params(self):
        """Re-derive the fp16 hash table and the packed MLP operands from the fp32 masters (after an external optimiser or
        a checkpoint load changed them; tcnn does this cast at every forward, tinycudann/modules.py:156)."""
        F.cast_f16(self.table, self.table_h)
        F.mlp_pack(self.geo_p, self.geo_spec, self.geo_blob)
        F.mlp_pack(self.rad_p, self.rad_spec, self.rad_blob)

    def set_occupancy(self, volume_bool):
        """(n,n,n) bool occupancy (numpy or tensor), the VolumeBound bitfield."""
        v = torch.as_tensor(volume_bool).to(self.dev, torch.bool).contiguous()
        assert v.shape == self.volume.shape
        self.volume.copy_(v)
        if self.use_bits:
            from . import synth
            self.bits.copy_(torch.from_numpy(synth.volume_to_bitfield(v.cpu().numpy())).to(self.dev))

    # ------------------------------------------------------------------------------------------ occupancy maintenance
    @torch.no_grad()
    def forward_pts(self, pts, chunk=1 << 19):
        """Density at arbitrary world points (N,3) -> (N,): hash encode + geometry head with the TruncExp epilogue, in
        chunks (the reference's Base3dModel.forward_pts through chunk_processing, base_3d_model.py:343-366)."""
        pts = pts.to(self.dev, torch.float32).contiguous()
        n = pts.shape[0]
        out = torch.empty((n, ), dtype=torch.float32, device=self.dev)
        m = min(n, chunk)
        if getattr(self, '_pts_buf', None) is None or self._pts_buf[0].shape[0] < m:
            self._pts_buf = (torch.empty((m, 3), dtype=torch.float32, device=self.dev),
                             torch.empty((m, self.enc_dim), dtype=torch.float16, device=self.dev),
                             torch.empty((m, self.cfg['w_feat']), dtype=torch.float16, device=self.dev))
        unit, enc, feat = self._pts_buf
        for a in range(0, n, chunk):
            b = min(n, a + chunk)
            _lib.call('arcb200_world_to_unit', pts[a:b].data_ptr(), self.aabb.data_ptr(), unit.data_ptr(), b - a, _lib.current_stream())
            F.grid_forward(unit[:b - a], self.table_h, self.geom, out=enc[:b - a])
            F.mlp_forward(enc[:b - a], self.geo_blob, self.geo_spec, out=feat[:b - a], exp0_out=out[a:b])
        return out

    @torch.no_grad()
    def forward_pts_dir(self, pts, dirs, chunk=1 << 19):
        """density and colour at world points seen along `dirs` (N,3 each) -> sigma (N,), rgb (N,3): one fused launch per chunk
        (Base3dModel.forward_pts_dir, base_3d_model.py:233-254; used for mesh colours, infer_func.py:582-632)"""
        pts = pts.to(self.dev, torch.float32).contiguous()
        dirs = dirs.to(self.dev, torch.float32).contiguous()
        n = pts.shape[0]
        sigma = torch.empty((n, ), dtype=torch.float32, device=self.dev)
        rgb = torch.empty((n, 3), dtype=torch.float32, device=self.dev)
        unit = torch.empty((min(n, chunk), 3), dtype=torch.float32, device=self.dev)
        nl, nf, off, sc, rs = self.geom._args()
        for a in range(0, n, chunk):
            b = min(n, a + chunk)
            _lib.call('arcb200_world_to_unit', pts[a:b].data_ptr(), self.aabb.data_ptr(), unit.data_ptr(), b - a, _lib.current_stream())
            _lib.call('arcb200_field_fwd', unit.data_ptr(), self.table_h.data_ptr(), nl, nf, off, sc, rs, self.geo_blob.data_ptr(),
                      self.rad_blob.data_ptr(), dirs[a:b].data_ptr(), None, 1, None, None, None, None, sigma[a:].data_ptr(),
                      rgb[a:].data_ptr(), None, b - a, None, _lib.current_stream())
        return sigma, rgb

    @torch.no_grad()
    def mesh_components(self, verts, faces):
        """get_mesh_components (infer_func.py:582-632): normals, face centres and the colours the model gives to vertices and
        face centres when looked at against their normals"""
        from .plugin.mesh import get_mesh_components
        vn, fn, fc = get_mesh_components(verts, faces)
        _, vc = self.forward_pts_dir(verts, -vn)
        _, fcol = self.forward_pts_dir(fc, -fn)
        return dict(verts=verts, faces=faces, vert_normals=vn, face_normals=fn, face_centers=fc, vert_colors=vc, face_colors=fcol)

    @torch.no_grad()
    def density_grid(self, n_grid, x_range=None, slab=16):
        """Density at the voxel centres of an n_grid^3 lattice over the bounding box (the grid a mesh extraction runs
        marching cubes on: Volume.get_volume_pts -> forward_pts, base_3d_model.py:343-366), for x slabs
        [x_range[0], x_range[1]) -> (nx, n_grid, n_grid) fp32.  Ranks take disjoint slabs (dist.shard_range) and gather."""
        from .plugin.mesh import volume_pts
        x0, x1 = x_range if x_range is not None else (0, n_grid)
        out = torch.empty((x1 - x0, n_grid, n_grid), dtype=torch.float32, device=self.dev)
        lo, hi = self.aabb[0].tolist(), self.aabb[1].tolist()
        pts = torch.empty((min(slab, max(1, x1 - x0)) * n_grid * n_grid, 3), dtype=torch.float32, device=self.dev)
        for a in range(x0, x1, slab):
            b = min(x1, a + slab)
            # voxel centres of the slab generated in place (csrc/mcubes.cu), the same values as Volume.get_volume_pts()
            p = volume_pts(lo, hi, n_grid, a, b, out=pts)
            out[a - x0:b - x0] = self.forward_pts(p).view(b - a, n_grid, n_grid)
        return out

    def extract_mesh(self, n_grid, level=50.0, grad_dir='descent', slab=16, x_range=None, return_grid=False):
        """run_infer_volume (arcnerf/eval/infer_func.py:448-500): density on the n_grid^3 lattice, then marching cubes
        (plugin/mesh.py) -> verts, faces, vertex normals as device tensors in world coordinates.
        x_range = (x0, x1): the cells of x slabs [x0, x1) only (a rank's shard, dist.shard_range); the slab reads lattice planes
        x0..x1, so neighbouring shards evaluate one plane twice and need no exchange."""
        from .plugin.mesh import extract_mesh
        x0, x1 = x_range if x_range is not None else (0, n_grid)
        sigma = self.density_grid(n_grid, x_range=(x0, min(n_grid, x1 + 1)), slab=slab)
        side = (self.aabb[1] - self.aabb[0]).tolist()
        lo = self.aabb[0].tolist()
        vs = [s / n_grid for s in side]
        if sigma.shape[0] < 2:
            verts = torch.zeros((0, 3), dtype=torch.float32, device=self.dev)
            faces, normals = torch.zeros((0, 3), dtype=torch.int32, device=self.dev), verts.clone()
        else:
            verts, faces, normals = extract_mesh(sigma, level, vs, side, grad_dir)
            # extract_mesh centres the volume at the origin; the reference adds volume_origin afterwards (infer_func.py:545-552)
            shift = torch.tensor([lo[0] + side[0] / 2 + x0 * vs[0], lo[1] + side[1] / 2, lo[2] + side[2] / 2],
                                 dtype=torch.float32, device=self.dev)
            verts = verts + shift
        return (verts, faces, normals, sigma) if return_grid else (verts, faces, normals)

    @torch.no_grad()
    def get_est_opacity(self, dt, pts):
        """base_3d_model.py:368-389: opacity = density * dt (the instant-ngp convention)"""
        out = self.forward_pts(pts)
        _lib.call('arcb200_scale_f32', out.data_ptr(), float(dt), out.numel(), _lib.current_stream())
        return out

    def attach_bound(self, bound):
        """Use an object bound (plugin/obj_bound.py VolumeBound) as the owner of the occupancy state; its bool volume is
        mirrored into the sampler's buffer after every refresh."""
        self.bound = bound.to(self.dev)
        if bound.volume.get_voxel_bitfield() is not None:
            self.volume.copy_(bound.volume.get_voxel_bitfield())

    @torch.no_grad()
    def update_occupancy(self, cur_epoch, apply=True):
        """FgModel.optimize (fg_model.py:400-410): obj_bound.optimize(cur_epoch, n_coarse_sample, get_est_opacity) every
        `epoch_optim` steps; warm-up sweeps every voxel, afterwards 1/4 random + 1/4 occupied voxels.
        apply=False does all the work but leaves the sampler's volume as it is (benchmarks on a fixed synthetic scene)."""
        bound = getattr(self, 'bound', None)
        if self.use_bits:
            # BitfieldBound: density grid + Morton bit field, the instant-ngp rule (bitfield_bound.py:130-189)
            if bound is None:
                from .plugin.obj_bound import BitfieldBound
                cfg = self.cfg
                bound = BitfieldBound(dict(bitfield=dict(n_grid=cfg['n_grid'], side=cfg['side']),
                                           epoch_optim=cfg.get('epoch_optim', 16),
                                           epoch_optim_warmup=cfg.get('epoch_optim_warmup', 256),
                                           near_distance=cfg['near_distance'], opa_thres=cfg.get('opa_thres', 0.01)))
                self.bound = bound = bound.to(self.dev)
                bound.density_bitfield.copy_(self.bits)
            bound.optimize(cur_epoch, self.cfg['n_sample'], self.get_est_opacity)
            if apply:
                self.bits.copy_(bound.density_bitfield)
                for ss in self.sets:
                    ss.key = None
            return
        if bound is None:
            from .plugin.obj_bound import VolumeBound
            cfg = self.cfg
            bound = VolumeBound(dict(volume=dict(n_grid=cfg['n_grid'], side=cfg['side']), epoch_optim=cfg.get('epoch_optim', 16),
                                     epoch_optim_warmup=cfg.get('epoch_optim_warmup', 256), ray_sample_acc=True,
                                     ray_sample_fix_step=True, near_distance=cfg['near_distance'],
                                     opa_thres=cfg.get('opa_thres', 0.01)))
            self.bound = bound.to(self.dev)
            self.bound.volume.bitfield.copy_(self.volume)          # the bound starts from the occupancy in use
            bound = self.bound
        bound.optimize(cur_epoch, self.cfg['n_sample'], self.get_est_opacity)
        if apply:
            self.volume.copy_(bound.volume.get_voxel_bitfield())
            for ss in self.sets:                # a prefetched sample set was marched through the old occupancy
                ss.key = None

    # ------------------------------------------------------------------------------------------ stages
    def _sample_into(self, ss, rays_o, rays_d):
        """slab test + packed occupancy march into sample set `ss` on the current stream; stays on the device"""
        cfg = self.cfg
        R = rays_o.shape[0]
        assert R <= self.max_rays
        st = _lib.current_stream()
        self._run('aabb', lambda: _lib.call(
            'arcb200_aabb_intersection', rays_o.data_ptr(), rays_d.data_ptr(), self.aabb.data_ptr(),
            ss.near.data_ptr(), ss.far.data_ptr(), None, ss.hit.data_ptr(), R, 1, st))
        self._run('march_packed', lambda: _lib.call(
            'arcb200_march_packed', int(self.use_bits), rays_o.data_ptr(), rays_d.data_ptr(), ss.near.data_ptr(),
            ss.far.data_ptr(), cfg['n_sample'], self.dt, self.aabb.data_ptr(), cfg['n_grid'],
            (self.bits if self.use_bits else self.volume).data_ptr(), cfg['near_distance'], R, self.rng.state, self.rng.inc,
            ss.scratch.data_ptr(), ss.counts.data_ptr(), ss.offsets.data_ptr(), ss.meta.data_ptr(),
            self.capacity, ss.t.data_ptr(), None, ss.ray_id.data_ptr(), ss.unit.data_ptr(), st))
        self.rng.advance()
        ss.key = (rays_o.data_ptr(), rays_d.data_ptr(), R)

    def sample(self, rays_o, rays_d):
        """Make the samples of this ray batch current: take the set a prefetch() filled for exactly these tensors (the
        main stream then only waits for its event), otherwise march now on the main stream."""
        key = (rays_o.data_ptr(), rays_d.data_ptr(), rays_o.shape[0])
        other = self.sets[1 - self.cur]
        if other.ready is not None and other.key == key:
            torch.cuda.current_stream().wait_event(other.ready)
            other.ready = None
        else:
            if other.ready is not None:                 # a prefetch for other rays is still writing this set
                torch.cuda.current_stream().wait_event(other.ready)
                other.ready = None
            self._sample_into(other, rays_o, rays_d)
        self.cur = 1 - self.cur
        self.forward_serial += 1

    def prefetch(self, rays_o, rays_d):
        """Sample the NEXT ray batch on the side stream while the current one is in the field kernels.  The set being
        overwritten was last read by the step before the current one; an event on the main stream orders that.
        The caller keeps rays_o / rays_d alive and unchanged until the batch has been consumed."""
        nxt = self.sets[1 - self.cur]
        gate = torch.cuda.Event()
        gate.record()                                  # everything enqueued so far on the main stream
        with torch.cuda.stream(self.side_stream):
            self.side_stream.wait_event(gate)
            self._sample_into(nxt, rays_o, rays_d)
            nxt.ready = torch.cuda.Event()
            nxt.ready.record()

    def field_forward(self, rays_d, clear_loss=False):
        """encode -> geometry head -> density | [feat, SH] -> radiance head, on the packed samples"""
        C, nd = self.capacity, self.n_dev
        cfg = self.cfg
        if self.fused_fwd:
            # one launch: encode -> geometry head (+ density) -> [feat | SH] -> radiance head (csrc/field_fused.cuh)
            tr = self.training
            nl, nf, off, sc, rs = self.geom._args()
            self._run('field_fwd', lambda: _lib.call(
                'arcb200_field_fwd', self.unit.data_ptr(), self.table_h.data_ptr(), nl, nf, off, sc, rs, self.geo_blob.data_ptr(),
                self.rad_blob.data_ptr(), rays_d.data_ptr(), self.ray_id.data_ptr(), 1, self.enc.data_ptr() if tr else None,
                self.rad_in.data_ptr() if tr else None, self.geo_act.data_ptr() if tr else None,
                self.rad_act.data_ptr() if tr else None, self.sigma.data_ptr(), self.rgb_s.data_ptr(),
                self.loss.data_ptr() if clear_loss else None, C, nd.data_ptr(), _lib.current_stream()))
            return
        self._run('grid_fwd', lambda: F.grid_forward(self.unit, self.table_h, self.geom, n_dev=nd, out=self.enc))
        # geometry head writes its W_feat outputs straight into the radiance input row; SH fills the rest
        # (the density sigma = TruncExp(channel 0) comes out of the same kernel's epilogue)
        tr = self.training
        fuse_sh = self.cfg['sh_degree'] == 4
        self._run('mlp_geo_fwd', lambda: F.mlp_forward(self.enc, self.geo_blob, self.geo_spec, out=self.rad_in, out_col=0,
                                                       n_dev=nd, act_save=self.geo_act if tr else None, exp0_out=self.sigma,
                                                       sh=(rays_d, self.ray_id, self.cfg['w_feat'], True, self.loss if clear_loss else None) if fuse_sh else None))
        if clear_loss and not fuse_sh:
            _lib.call('arcb200_fill_f32', self.loss.data_ptr(), 0.0, 1, _lib.current_stream())
        if not fuse_sh:
            self._run('sh_fwd', lambda: F.sh_forward(rays_d, self.cfg['sh_degree'], index=self.ray_id, remap=True, out=self.rad_in,
                                                     out_col=self.cfg['w_feat'], n_dev=nd))
        self._run('mlp_rad_fwd', lambda: F.mlp_forward(self.rad_in, self.rad_blob, self.rad_spec, out=self.rgb_s, n_dev=nd,
                                                       act_save=self.rad_act if tr else None))

    def composite(self, R, bkg):
        bkg_per_ray = int(bkg is not None and bkg.shape[0] == R and R > 1)
        self._run('composite_fwd', lambda: _lib.call(
            'arcb200_composite_fwd_ex', self.offsets.data_ptr(), 0, self.capacity, self.pmax.data_ptr(),
            self.sigma.data_ptr(), 0, self.rgb_s.data_ptr(), self.t.data_ptr(), None, 0, 0, _lib.ptr(bkg),
            bkg_per_ray, self.rgb.data_ptr(), self.depth.data_ptr(), self.acc.data_ptr(), None, None, None,
            self.t_last.data_ptr(), R, _lib.current_stream()))

    def forward(self, rays_o, rays_d, bkg=None, next_rays=None):
        """-> rgb (R,3), depth (R,), mask (R,) views of the pipeline's output buffers.
        next_rays = (rays_o, rays_d) of the batch that follows: its march is started on the side stream."""
        R = rays_o.shape[0]
        self.sample(rays_o, rays_d)
        if next_rays is not None:
            self.prefetch(*next_rays)
        self.field_forward(rays_d)
        self.composite(R, bkg)
        return self.rgb[:R], self.depth[:R], self.acc[:R]

    def backward(self, rays_d, R, bkg, composite=True, g_depth=None, g_mask=None):
        """backward of forward() from self.g_rgb (composite=False: d_sigma / d_rgb_s are already there)"""
        cfg, C, nd, st = self.cfg, self.capacity, self.n_dev, _lib.current_stream()
        bkg_per_ray = int(bkg is not None and bkg.shape[0] == R and R > 1)
        if composite:
            self._run('composite_bwd', lambda: _lib.call(
                'arcb200_composite_bwd', self.offsets.data_ptr(), 0, C, self.pmax.data_ptr(), self.sigma.data_ptr(), 0,
                self.rgb_s.data_ptr(), self.t.data_ptr(), None, 0, 0, _lib.ptr(bkg), bkg_per_ray,
                self.g_rgb.data_ptr(), _lib.ptr(g_depth), _lib.ptr(g_mask), None, self.d_sigma.data_ptr(),
                self.d_rgb_s.data_ptr(), cfg['n_sample'], R, st))
        self._run('mlp_rad_bwd', lambda: F.mlp_backward(self.rad_in, self.rad_blob, self.rad_spec, self.d_rgb_s,
                                                        dW=self.g_rad, dx=self.d_rad_in, n_dev=nd,
                                                        act_saved=self.rad_act, y_saved=self.rgb_s))
        # geometry head: feature gradient = first W_feat columns of d_rad_in; the density gradient joins channel 0
        # through TruncExp inside the kernel (no separate merge pass)
        self._run('mlp_geo_bwd', lambda: F.mlp_backward(self.enc, self.geo_blob, self.geo_spec, self.d_rad_in,
                                                        dW=self.g_geo, dx=self.d_enc, n_dev=nd, act_saved=self.geo_act,
                                                        g0=self.d_sigma, x0=self.rad_in, clip=cfg['trunc_exp_clip']))
        self._run('grid_bwd', lambda: F.grid_backward(self.unit, self.d_enc, self.geom, grad_table=self.g_table, n_dev=nd))

    def optimizer_step(self, counted=False):
        """counted: the device step counter was already advanced by this step's compositing launch"""
        cfg = self.cfg
        self.step_count += 1
        gs = 1.0 / self.world_size
        if not counted:
            _lib.call('arcb200_counter_add', self.step_dev.data_ptr(), 1, _lib.current_stream())
        kw = dict(beta1=cfg['beta1'], beta2=cfg['beta2'], eps=cfg['eps'], weight_decay=cfg['weight_decay'], grad_scale=gs,
                  step_dev=self.step_dev)
        if getattr(self, '_mlp_opt', None) is None:
            self._mlp_opt = F.MlpAdamPack((self.geo_spec, self.rad_spec), (self.geo_p, self.rad_p), (self.m_geo, self.m_rad),
                                          (self.v_geo, self.v_rad), (self.g_geo, self.g_rad), (self.geo_blob, self.rad_blob))

        def mlp_update():
            self._mlp_opt.step(self.step_count, cfg['lr'], **kw)

        def table_update():
            F.adam_step(self.table, self.m_table, self.v_table, self.g_table, self.step_count, cfg['lr'], p_half=self.table_h, **kw)
        self._run('adam_table', table_update)
        self._run('adam_mlp+pack', mlp_update)

    def train_step(self, rays_o, rays_d, target, bkg=None, optimize=True, next_rays=None):
        """One optimisation step on a ray batch; returns the (device) loss tensor.  No host synchronisation.
        next_rays = (rays_o, rays_d) of the following batch, if known: it is sampled on the side stream meanwhile
        (the sampler does not depend on the parameters being updated)."""
        R = rays_o.shape[0]
        self.sample(rays_o, rays_d)
        if next_rays is not None:
            self.prefetch(*next_rays)
        self._step_core(R, rays_d, target, bkg, count_step=optimize)
        if optimize:
            if self.world_size > 1:
                self.exchange_and_step(counted=True)
            else:
                self.optimizer_step(counted=True)
        return self.loss

    def exchange_and_step(self, counted=False):
        """Data-parallel step on the mean gradient over ranks - what DDP + Adam deliver in the reference
        (common/trainer/basic_trainer.py:192-202).
        'peer': barrier, then one kernel per parameter vector that sums the peers' gradients over NVLink, applies Adam
        and broadcasts the fp16 working copy (hash table: this rank's slice only; the two small heads: redundantly on
        every rank, same summation order, so the replicas stay identical), barrier, clear the own gradients.
        'nccl': ONE all-reduce of the flat gradient buffer (48.8 MB), then the single-GPU Adam kernels."""
        if self.exchange not in ('peer', 'peer16', 'mc16'):
            import torch.distributed as dist
            self._run('all_reduce', lambda: dist.all_reduce(self.g_all))
            self.optimizer_step(counted)
            return
        from .peer import adam_step_peer
        cfg, ar, W = self.cfg, self.arena, self.world_size
        self.step_count += 1
        if not counted:
            _lib.call('arcb200_counter_add', self.step_dev.data_ptr(), 1, _lib.current_stream())
        kw = dict(beta1=cfg['beta1'], beta2=cfg['beta2'], eps=cfg['eps'], weight_decay=cfg['weight_decay'],
                  grad_scale=1.0 / W, step_dev=self.step_dev)
        nt, ng = self.table.numel(), self.geo_p.numel()
        if not hasattr(self, '_peer_ptrs'):
            self._peer_ptrs = (ar.pointers('grad'), ar.pointers('grad', nt * 4), ar.pointers('grad', (nt + ng) * 4),
                               ar.pointers('table_h'))
        g_t, g_g, g_r, th = self._peer_ptrs
        lo, hi = self.slice
        half = self.exchange in ('peer16', 'mc16')
        mc = self.exchange == 'mc16'
        if half:
            from .peer import adam_step_mc16, adam_step_peer16, grad_to_half
            if not hasattr(self, '_g16'):
                self._g16 = self.arena.local('grad16', torch.float16)
                self._g16_ptrs = ar.pointers('grad16')
            # 'mc16': the whole vector goes to fp16 (the switch sums every rank's copy, the own one included)
            keep = (0, 0) if mc else (lo, hi)
            self._run('grad_to_half', lambda: grad_to_half(self.g_table, self._g16, keep[0], keep[1], cfg.get('exchange_scale', 128.0)))
        self._run('peer_barrier', ar.barrier)
        def table_update():
            if mc:
                adam_step_mc16(ar.mc_pointer('grad16'), ar.mc_pointer('table_h'), lo, hi, self.table, self.m_table, self.v_table,
                               self.step_count, cfg['lr'], scale=cfg.get('exchange_scale', 128.0), **kw)
            elif half:
                adam_step_peer16(self.g_table, self._g16_ptrs, th, W, self.rank, lo, hi, self.table, self.m_table, self.v_table,
                                 self.step_count, cfg['lr'], scale=cfg.get('exchange_scale', 128.0), **kw)
            else:
                adam_step_peer(g_t, th, W, lo, hi, self.table, self.m_table, self.v_table, self.step_count, cfg['lr'], **kw)

        if getattr(self, '_mlp_opt_peer', None) is None:
            self._mlp_opt_peer = F.MlpAdamPack((self.geo_spec, self.rad_spec), (self.geo_p, self.rad_p),
                                               (self.m_geo, self.m_rad), (self.v_geo, self.v_rad),
                                               ([x for x in g_g], [x for x in g_r]), (self.geo_blob, self.rad_blob))

        def mlp_update():
            self._mlp_opt_peer.step(self.step_count, cfg['lr'], zero_grad=False, **kw)
        if self.timer is not None or not self.overlap_heads:
            self._run('adam_table_peer', table_update)
            self._run('adam_mlp+pack', mlp_update)
        else:
            # the table kernel waits on NVLink, not on SMs: the heads' update (every rank's 40 KB of gradients, one small launch)
            # runs beside it on a second stream and is joined before the closing barrier
            main = torch.cuda.current_stream()
            self._opt_fork.record(main)
            self.opt_stream.wait_event(self._opt_fork)
            with torch.cuda.stream(self.opt_stream):
                mlp_update()
                self._opt_join.record(self.opt_stream)
            table_update()
            main.wait_event(self._opt_join)

        def close():
            ar.barrier()
            if half:
                heads = self.g_all[nt:]          # the heads' gradients (40 KB); the table part was cleared on the way
                _lib.call('arcb200_fill_f32', heads.data_ptr(), 0.0, heads.numel(), _lib.current_stream())
            else:
                self.g_all.zero_()
        self._run('peer_barrier+clear', close)

    def close(self):
        """Orderly teardown of a multi-rank pipeline: captured graphs first (they reference arena memory), then the peer arena
        (every rank unmaps its peers, a barrier, every rank frees its own allocation).  Collective: call it on every rank, before
        torch.distributed.destroy_process_group()."""
        self.graphs = [None, None]
        torch.cuda.synchronize(self.dev)
        if self.arena is not None:
            import torch.distributed as dist
            arena, self.arena = self.arena, None
            arena.close(sync_ranks=dist.barrier)

    def gather_master_table(self):
        """'peer' exchange keeps the fp32 master of the hash table sharded (each rank owns self.slice); this collects
        the full fp32 table on every rank (checkpointing; not on the step path)."""
        if self.exchange not in ('peer', 'peer16', 'mc16'):
            return self.table
        import torch.distributed as dist
        from .peer import slice_bounds
        nt = self.table.numel()
        per = slice_bounds(nt, 0, self.world_size)[1]
        pad = torch.zeros(per, dtype=torch.float32, device=self.dev)
        lo, hi = self.slice
        pad[:hi - lo] = self.table[lo:hi]
        if dist.get_backend() == 'gloo':            # gloo gathers host tensors only
            pad = pad.cpu()
        outs = [torch.empty_like(pad) for _ in range(self.world_size)]
        dist.all_gather(outs, pad)
        self.table.copy_(torch.cat(outs)[:nt])
        if self.arena is not None:
            self.arena.check()                      # the host is synchronised here anyway: a timed-out barrier is reported
        return self.table

    # ------------------------------------------------------------------------------------------ replayed step
    def _step_core(self, R, rays_d, target, bkg, count_step=True):
        """field forward -> compositing + Huber loss + compositing backward (one kernel) -> field backward, on the current
        sample set; fills self.rgb, self.g_rgb, self.loss and the flat gradient buffer.  The loss accumulator is cleared by
        the geometry head's launch and the device step counter advanced by the compositing launch (no launches of their own)."""
        self.field_forward(rays_d, clear_loss=True)
        st = _lib.current_stream()
        bkg_per_ray = int(bkg is not None and bkg.shape[0] == R and R > 1)
        self._run('composite_train', lambda: _lib.call(
            'arcb200_composite_huber_train', self.offsets.data_ptr(), 0, self.capacity, self.pmax.data_ptr(),
            self.sigma.data_ptr(), self.rgb_s.data_ptr(), self.t.data_ptr(), None, 0, 0, _lib.ptr(bkg), bkg_per_ray,
            target.data_ptr(), self.cfg['huber_delta'], self.cfg['loss_weight'], self.rgb.data_ptr(), self.g_rgb.data_ptr(),
            self.loss.data_ptr(), self.d_sigma.data_ptr(), self.d_rgb_s.data_ptr(),
            self.step_dev.data_ptr() if count_step else None, self.cfg['n_sample'], R, st))
        self.backward(rays_d, R, bkg, composite=False)

    def _step_body(self, R, rays_d, target, bkg):
        """everything of a training step after the sampler, on the current sample set and the current stream"""
        self._step_core(R, rays_d, target, bkg)
        if self.world_size > 1:
            self.exchange_and_step(counted=True)
        else:
            self.optimizer_step(counted=True)

    def enable_graphs(self, n_rays):
        """Static input slots + one CUDA graph per sample-set parity for the post-sampler part of a step (20 launches
        replayed as one).  The sampler itself stays a normal launch on the side stream: its jitter generator state is
        a by-value argument that changes every call.  With several ranks the gradient all-reduce is captured too."""
        dev, f32 = self.dev, torch.float32
        self.static_in = [tuple(torch.empty((n_rays, 3), dtype=f32, device=dev) for _ in range(4)) for _ in range(2)]
        self.graph_rays = n_rays
        self.graphs = [None, None]
        self._gate_events = [torch.cuda.Event(), torch.cuda.Event()]
        self._ready_events = [torch.cuda.Event(), torch.cuda.Event()]
        rng_state = (self.rng.state, self.rng.inc)       # the warm-up marches must not consume the jitter stream
        # warm up outside capture (lazy module loading, function attributes), on both parities
        for _ in range(2):
            for x in self.static_in[1 - self.cur]:
                x.zero_()
            o, d, t, b = self.static_in[1 - self.cur]
            d[:, 2] = 1.0
            self.train_step(o, d, t, b, optimize=False)
        if self.exchange == 'nccl':                      # communicator set-up must not happen inside the capture
            import torch.distributed as dist
            dist.all_reduce(self.g_all)
        torch.cuda.synchronize()
        self.g_all.zero_()
        count0, step0 = self.step_count, self.step_dev.clone()
        keep = (self.table.clone(), self.geo_p.clone(), self.rad_p.clone(), self.m_table.clone(), self.v_table.clone(),
                self.m_geo.clone(), self.v_geo.clone(), self.m_rad.clone(), self.v_rad.clone())
        saved_cur = self.cur
        for p in range(2):
            self.cur = p
            o, d, t, b = self.static_in[p]
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._step_body(n_rays, d, t, b)
            self.graphs[p] = g
        self.cur = saved_cur
        # capture does not execute, but restore the optimiser state anyway in case of eager fall-backs above
        for dst, src in zip((self.table, self.geo_p, self.rad_p, self.m_table, self.v_table, self.m_geo, self.v_geo,
                             self.m_rad, self.v_rad), keep):
            dst.copy_(src)
        self.step_count = count0
        self.step_dev.copy_(step0)
        self.rng.state, self.rng.inc = rng_state
        for ss in self.sets:
            ss.ready, ss.key = None, None
        self.table_h.copy_(self.table.half())
        F.mlp_pack(self.geo_p, self.geo_spec, self.geo_blob)
        F.mlp_pack(self.rad_p, self.rad_spec, self.rad_blob)
        torch.cuda.synchronize()

    def _upload_and_sample(self, slot, batch):
        """batch (rays_o, rays_d, target, bkg) host-pinned or device tensors -> static slot, then march (current stream)"""
        for dst, src in zip(self.static_in[slot], batch):
            dst.copy_(src, non_blocking=True)
        ss = self.sets[slot]
        self._sample_into(ss, self.static_in[slot][0], self.static_in[slot][1])

    def train_step_graphed(self, batch, next_batch=None):
        """One optimisation step from a (host or device) batch through the static slots and the captured graph.
        next_batch, if given, is uploaded and sampled on the side stream while this step runs.  Returns the device
        loss tensor."""
        p = 1 - self.cur
        ss = self.sets[p]
        main = torch.cuda.current_stream()
        if ss.ready is not None and ss.key == ('slot', p, id(batch[0])):
            main.wait_event(ss.ready)
        else:
            if ss.ready is not None:
                main.wait_event(ss.ready)
            self._upload_and_sample(p, batch)
        ss.ready = None
        self.cur = p
        self.forward_serial += 1
        if next_batch is not None:
            gate = self._gate_events[p]
            gate.record()                      # the side stream may touch slot 1-p once everything queued so far is done
        self.step_count += 1
        self.graphs[p].replay()                # the GPU starts on this step before the CPU prepares the next one
        if next_batch is not None:
            q = 1 - p
            with torch.cuda.stream(self.side_stream):
                self.side_stream.wait_event(gate)
                self._upload_and_sample(q, next_batch)
                self.sets[q].key = ('slot', q, id(next_batch[0]))
                self.sets[q].ready = self._ready_events[q]
                self.sets[q].ready.record()
        return self.loss

    @torch.no_grad()
    def render(self, rays_o, rays_d, bkg=None, chunk=None):
        """Full-frame forward in chunks of `max_rays` rays (the reference chunks by chunk_rays=32768,
        full_model.py:415-457); returns rgb (N,3), depth (N,), mask (N,)."""
        was_training, self.training = self.training, False      # inference: no activations kept for a backward
        try:
            return self._render(rays_o, rays_d, bkg, chunk)
        finally:
            self.training = was_training

    def _render(self, rays_o, rays_d, bkg, chunk):
        N = rays_o.shape[0]
        chunk = chunk or self.max_rays
        rgb = torch.empty((N, 3), dtype=torch.float32, device=self.dev)
        depth = torch.empty((N,), dtype=torch.float32, device=self.dev)
        acc = torch.empty((N,), dtype=torch.float32, device=self.dev)
        for s in range(0, N, chunk):
            e = min(N, s + chunk)
            b = None if bkg is None else (bkg if bkg.shape[0] == 1 else bkg[s:e])
            nxt = (rays_o[e:min(N, e + chunk)], rays_d[e:min(N, e + chunk)]) if e < N else None
            r, d, a = self.forward(rays_o[s:e], rays_d[s:e], b, next_rays=nxt)
            rgb[s:e].copy_(r); depth[s:e].copy_(d); acc[s:e].copy_(a)
        return rgb, depth, acc

    # number of library kernels one train_step enqueues (for bench.py's gpu_launches)
    LAUNCHES_FORWARD = 1 + 4 + 1 + 1                  # aabb, march(4), fused field forward (encode + both heads), composite
    LAUNCHES_TRAIN = LAUNCHES_FORWARD + 1 + 1 + 1 + 1 + 1   # (composite fused with loss + its backward + step count), rad bwd, geo bwd, grid bwd, table adam, heads adam+pack
    LAUNCHES_REFRESH = 4 + 2 * 3 + 4                   # candidates (count, scan, select, points), 2 chunks x (unit, grid, geo mlp), apply (fill, splat, ema, threshold)
    LAUNCHES_TRAIN_PEER = LAUNCHES_TRAIN + 2           # + two flag barriers (the gradient clear is a memset, not a kernel of ours)
