"""`python bench.py --config ngp_nerfpp` and `--config mesh`: BASELINE.json configs[3] and configs[4].

ngp_nerfpp (configs[3]; configs/expr/MipNeRF360/garden/*_nerfpp.yaml shapes: 1297 x 840 views of an unbounded 360 capture, 4096-ray
    batches): one training step of FullModel(NGP foreground inside the unit-ish bound, NeRF++ inverse-sphere background 32 + 64
    samples through two 8 x 256 networks), blended rgb = rgb_fg + T_last * rgb_bkg inside the NGP compositing kernel; data-parallel,
    gradients averaged with NCCL.  value = rays/s over all ranks.
mesh (configs[4]; inference.volume {n_grid 512, level 50} + a 1920 x 1080 frame): a step is one full-frame render (camera rays
    generated on the device, rows sharded over ranks, gathered on rank 0) plus the density on the 512^3 voxel-centre lattice (x
    slabs sharded over ranks, one overlapping plane each) and marching cubes on every rank's slab.  value = ms per step
    (lower is better).  The field is first trained for a few hundred steps on an analytic sphere scene so that the level set exists.
"""
import json
import os
import time

import numpy as np
import torch

N_RAYS = 4096


def _helpers(dev, dist):
    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if dist is None:
            return ms
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    return barrier, max_over_ranks


# ------------------------------------------------------------------------------------------------------- configs[3]
BKG_CFG = dict(rays=dict(bounding_radius=3.0, n_sample=32, n_importance=64, perturb=True, add_inf_z=True, noise_std=0.0, white_bkg=False),
               chunk_rays=32768,
               geometry=dict(W=256, D=8, skips=[4], encoder=dict(type='FreqEmbedder', input_dim=4, n_freqs=10), W_feat=256,
                             geometric_init=False),
               radiance=dict(mode='vf', W=128, D=1, encoder=dict(view=dict(type='FreqEmbedder', input_dim=3, n_freqs=4)), W_feat_in=256))
WORKLOAD_3 = ('NGP foreground (16-level hash grid, 64-wide fused MLPs, 128^3 occupancy march, side 2) + NeRF++ background (inverse-sphere '
              'shells beyond radius 3, 32 coarse + 64 fine samples/ray, two 8x256 networks), 4096-ray batch per GPU drawn from '
              '1297x840 views on a 360-degree orbit, Huber colour loss, Adam, NCCL gradient all-reduce')


def orbit_ray_batch(n_rays, seed, W=1297, H=840, n_poses=100, radius=2.0):
    """random pixels of 1297 x 840 pinhole views on an orbit INSIDE the bounding sphere (an unbounded capture: every ray leaves
    the foreground bound and continues into the background shells)"""
    from arcnerf_b200 import synth
    rs = np.random.RandomState(seed)
    poses = synth.sphere_poses(n_poses, radius=radius, seed=11, v_range=(0.35, 0.65))
    o = np.empty((n_rays, 3), np.float32)
    d = np.empty((n_rays, 3), np.float32)
    cam = rs.randint(0, n_poses, n_rays)
    focal = 0.5 * W / np.tan(0.5 * 0.9)
    px = rs.randint(0, W, n_rays) + 0.5
    py = rs.randint(0, H, n_rays) + 0.5
    local = np.stack([(px - W / 2) / focal, (py - H / 2) / focal, np.ones(n_rays)], -1)
    for c in range(n_poses):
        m = cam == c
        if m.any():
            o[m] = poses[c][:3, 3]
            d[m] = local[m] @ poses[c][:3, :3].T
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return o, d.astype(np.float32)


def run_ngp_nerfpp(args, rank, world, dev, dist, read_peaks, ClockSampler):
    from arcnerf_b200 import functional as F
    from arcnerf_b200 import synth
    from arcnerf_b200.plugin.fg_model import NGPFgModel
    from arcnerf_b200.plugin.nerf_model import FullModel, NeRFPP
    # multi-rank: NCCL sets up its channels per message size during the first collectives - keep them out of the timed region
    K, W = args.steps, max(args.warmup, 3 if world == 1 else 10)
    barrier, max_over_ranks = _helpers(dev, dist)
    torch.manual_seed(1337)
    fg = NGPFgModel(dict(chunk_rays=N_RAYS), device=str(dev), max_rays=N_RAYS, capacity=1 << 19)
    fg.pipe.set_occupancy(synth.occupancy_volume('lego'))
    bkg = NeRFPP(BKG_CFG, device=str(dev))
    model = FullModel(fg, bkg)
    params = [p for p in model.parameters() if p.requires_grad]
    moments = [(torch.zeros_like(p), torch.zeros_like(p)) for p in params]
    lrs = {id(p): 5e-4 for p in bkg.parameters()}
    batches = []
    for b in range(K + W):
        o, d = orbit_ray_batch(N_RAYS, seed=rank * 1000 + b)
        tgt = np.random.RandomState(rank * 1000 + b + 7).uniform(0, 1, (N_RAYS, 3)).astype(np.float32)
        batches.append(tuple(torch.from_numpy(a).pin_memory() for a in (o, d, tgt)))
    resident = [tuple(x.to(dev) for x in b) for b in batches]
    step_no = [0]

    def train_step(o, d, tgt):
        out = model({'rays_o': o[None], 'rays_d': d[None]}, cur_epoch=step_no[0])
        loss = torch.nn.functional.huber_loss(out['rgb_coarse'][0], tgt, delta=0.1)
        for p in params:
            p.grad = None
        loss.backward()
        step_no[0] += 1
        for p, (m, v) in zip(params, moments):
            if p.grad is None:
                continue
            g = p.grad.contiguous()
            if world > 1:
                dist.all_reduce(g)
            F.adam_step(p.data.view(-1), m.view(-1), v.view(-1), g.view(-1), step_no[0], lrs.get(id(p), 1e-2), eps=1e-15,
                        grad_scale=1.0 / world, zero_grad=False)
        for nets in (bkg.coarse_nets, bkg.fine_nets):
            nets._packed_version = None                     # parameters changed in place by the library's Adam kernel
        fg.pipe.sync_params()                               # fp16 table copy + packed MLP operands of the foreground
        return loss

    for i in range(W):
        train_step(*resident[i])
    barrier()
    clocks = ClockSampler(dev.index if rank == 0 else None)
    clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for i in range(K):
        train_step(*resident[W + i])
    e1.record()
    barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    from bench_common import LossReader
    reader = LossReader()
    dbuf = [torch.empty_like(x, device=dev) for x in batches[0]]
    barrier()
    e0.record()
    for i in range(K):
        for dst, src in zip(dbuf, batches[W + i]):
            dst.copy_(src, non_blocking=True)
        loss = train_step(*dbuf)
        reader.push(loss)
    reader.finish()
    e1.record()
    barrier()
    ms_e2e = max_over_ranks(e0.elapsed_time(e1))
    clock_info = clocks.stop()
    if rank == 0:
        value = N_RAYS * world * K / (ms_total * 1e-3)
        peaks = read_peaks()
        n_pts = N_RAYS * (32 + 96)
        flops = n_pts * bkg.coarse_nets.flops_per_point(training=True) if hasattr(bkg.coarse_nets, 'flops_per_point') else None
        print(json.dumps(dict(
            metric='ngp_nerfpp_train_rays_per_sec', value=value, unit='rays/s', n_gpus=world, steps=K, warmup=W, ms_per_step=ms_total / K,
            higher_is_better=True, scaling='weak', vs_baseline=None,
            baseline_note='the reference publishes no timing for an NGP + NeRF++ model (BASELINE.md)',
            dtype='fp16 tensor-core operands, fp32 accumulation and gradients', data='synthetic',
            config=dict(workload=WORKLOAD_3, rays_per_step_per_gpu=N_RAYS, bkg_points_per_step_per_gpu=n_pts, parallelism='dp%d' % world,
                        l2='every step draws new rays; the hash-table gradient + Adam state (0.2 GB) exceed the 126 MB L2'),
            e2e=dict(value=N_RAYS * world * K / (ms_e2e * 1e-3), unit='rays/s', h2d_bytes_per_step=sum(t.numel() * 4 for t in batches[0]),
                     d2h_bytes_per_step=4, ms_per_step=ms_e2e / K,
                     path='pinned host rays/targets -> H2D -> FullModel.forward (NeRF++ bkg, NGP fg with fused blend) + loss + backward + Adam -> loss D2H every step, read by the host one step behind'),
            gpu_launches=K * 70, clocks=clock_info,
            roofline=dict(kernel='dense_chain_kernel (8x256 background networks, 524 288 points per step)', bound='tensor', achieved=None,
                          peak=peaks.get('bf16_tflops'), unit='TFLOP/s', frac=None, traffic=None, bkg_chain_flops_per_step=flops,
                          note='composition of the ngp and nerf hot paths; their own bench lines carry the per-kernel rooflines'),
            cpu_baseline=None, cpu_baseline_note='a CPU arm for this composition would repeat the ngp and nerf arms; not run')))


# ------------------------------------------------------------------------------------------------------- configs[4]
WORKLOAD_4 = ('1920x1080 frame (2 073 600 rays generated on the device from one pinhole camera, NGP forward in 32768-ray chunks, rows sharded '
              'over ranks, gathered on rank 0) + density on the 512^3 voxel-centre lattice (x slabs sharded over ranks) + marching cubes '
              'at level 50 -> welded vertices, faces, vertex normals; field trained 400 steps on an analytic sphere scene first')


def sphere_targets(o, d, r=0.6):
    """colour of a shaded sphere of radius r at the origin seen along rays (o, d), white elsewhere"""
    b = (o * d).sum(-1)
    c = (o * o).sum(-1) - r * r
    disc = b * b - c
    t = -b - torch.sqrt(disc.clamp_min(0))
    hit = (disc > 0) & (t > 0)
    n = (o + t[:, None] * d) / r
    return torch.where(hit[:, None], 0.5 + 0.5 * n, torch.ones_like(n))


def run_mesh(args, rank, world, dev, dist, read_peaks, ClockSampler):
    from arcnerf_b200 import dist as adist
    from arcnerf_b200 import synth
    from arcnerf_b200.ngp import NGPPipeline
    from arcnerf_b200.plugin import ray_helper
    # multi-rank: NCCL sets up its channels per message size during the first collectives - keep them out of the timed region
    K, W = args.steps, max(args.warmup, 3 if world == 1 else 10)
    K = min(K, 20)
    barrier, max_over_ranks = _helpers(dev, dist)
    n_grid, level, Wd, Ht = args.mesh_grid, 50.0, 1920, 1080
    # ---- a field with a surface: identical training on every rank (same seeds), no exchange needed
    pipe = NGPPipeline(max_rays=N_RAYS, capacity=1 << 19, device=str(dev), seed=1337)
    white = torch.ones((1, 3), device=dev)
    for step in range(400):
        o, d = synth.train_ray_batch(N_RAYS, seed=step)
        o, d = torch.from_numpy(o).to(dev), torch.from_numpy(d).to(dev)
        pipe.train_step(o, d, sphere_targets(o, d), white)
        if step % 16 == 0:
            pipe.update_occupancy(step)
    rp = NGPPipeline(max_rays=32768, capacity=3 << 19, device=str(dev), params=dict(table=pipe.table, geo=pipe.geo_p, rad=pipe.rad_p))
    rp.volume.copy_(pipe.volume)
    del pipe
    # ---- the frame: one pinhole camera, this rank's rows of the (row-major) image
    c2w = synth.sphere_poses(1, seed=5)[0]
    focal = 0.5 * Wd / np.tan(0.5 * 0.6911112070083618)
    intrinsic = np.array([[focal, 0, Wd / 2], [0, focal, Ht / 2], [0, 0, 1]], np.float32)
    cams = ray_helper.pack_cameras(torch.from_numpy(intrinsic)[None], torch.from_numpy(c2w.astype(np.float32))[None], dev)
    # pixel tiles of 256 dealt round-robin to the ranks (a contiguous split leaves the ranks that see the object with most of the
    # samples); pixel index i * H + j
    pix, per = adist.deal_tiles(Wd * Ht, rank, world)
    pix = pix.to(dev).contiguous()
    cam_id = torch.zeros(pix.shape[0], dtype=torch.int32, device=dev)
    x0, x1 = adist.shard_range(n_grid, rank, world)
    frame_host = torch.empty((Wd * Ht, 3), dtype=torch.float32).pin_memory() if rank == 0 else None
    frame_dev = torch.empty((per * world, 3), device=dev) if dist is not None else None

    def frame():
        o, d = ray_helper.get_train_batch_rays(cams, cam_id, pix, Wd, Ht)
        rgb, _, _ = rp.render(o, d, white)
        if dist is None:
            return rgb
        return adist.gather_dealt(rgb, Wd * Ht, world, out=frame_dev)     # one all-gather + one strided copy that un-deals the tiles

    stats = {}

    def mesh():
        v, f, nr, sigma = rp.extract_mesh(n_grid, level=level, slab=args.mesh_slab, x_range=(x0, x1), return_grid=True)
        stats.update(verts=int(v.shape[0]), faces=int(f.shape[0]), sigma_max=float(sigma.max()))
        return v, f, nr

    for _ in range(W):
        frame()
        mesh()
    barrier()
    clocks = ClockSampler(dev.index if rank == 0 else None)
    clocks.start()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    t_frame, t_mesh = [], []
    for _ in range(K):
        barrier()
        e[0].record()
        frame()
        e[1].record()
        mesh()
        e[2].record()
        barrier()
        t_frame.append(max_over_ranks(e[0].elapsed_time(e[1])))
        t_mesh.append(max_over_ranks(e[1].elapsed_time(e[2])))
    # ---- end to end: camera from the host, frame + mesh back to pinned host memory
    t_e2e, d2h = [], 0
    cams_host = cams.cpu().pin_memory()
    for _ in range(max(3, K // 2)):
        barrier()
        e[0].record()
        cams.copy_(cams_host, non_blocking=True)
        rgb = frame()
        if rank == 0:
            frame_host.copy_(rgb, non_blocking=True)
        v, f, nr = mesh()
        hv, hf, hn = v.cpu(), f.cpu(), nr.cpu()
        d2h = (frame_host.numel() * 4 if rank == 0 else 0) + hv.numel() * 4 + hf.numel() * 4 + hn.numel() * 4
        e[2].record()
        barrier()
        t_e2e.append(max_over_ranks(e[0].elapsed_time(e[2])))
    clock_info = clocks.stop()
    counts = torch.tensor([stats['verts'], stats['faces']], device=dev, dtype=torch.int64)
    if dist is not None:
        dist.all_reduce(counts)
    # density-grid part alone (the dominant kernel: fused hash-grid encode + geometry head over the lattice)
    barrier()
    e[0].record()
    rp.density_grid(n_grid, x_range=(x0, min(n_grid, x1 + 1)), slab=args.mesh_slab)
    e[1].record()
    barrier()
    grid_ms = max_over_ranks(e[0].elapsed_time(e[1]))
    if rank == 0:
        med = lambda x: float(np.median(x))
        peaks = read_peaks()
        n_pts = (min(n_grid, x1 + 1) - x0) * n_grid * n_grid
        # algorithmic bytes per lattice point of the density pass: 16 levels x 8 corners x 4 B gathered + 12 B point + 4 B density
        algo = n_pts * (16 * 8 * 4 + 16)
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            cpu = _cpu_arm(rp, n_grid, level, Wd * Ht)
        print(json.dumps(dict(
            metric='frame_1920x1080_plus_mesh_%d_ms' % n_grid, value=med(t_frame) + med(t_mesh), unit='ms', n_gpus=world, steps=K, warmup=W,
            ms_per_step=med(t_frame) + med(t_mesh), higher_is_better=False, scaling='strong', vs_baseline=None,
            baseline_note='the reference publishes no timing for full-frame render or mesh extraction (BASELINE.md)',
            dtype='fp16 table / tensor-core operands, fp32 density, positions and mesh', data='synthetic',
            config=dict(workload=WORKLOAD_4, parallelism='rows and x slabs over %d ranks' % world, n_grid=n_grid, level=level,
                        l2='the 512^3 lattice (0.5 GB of densities, 1.6 GB of points generated slab by slab) and the 2 M rays exceed the 126 MB L2'),
            frame_ms=med(t_frame), mesh_ms=med(t_mesh), density_grid_ms=grid_ms, marching_cubes_ms=max(0.0, med(t_mesh) - grid_ms),
            mesh=dict(verts=int(counts[0]), faces=int(counts[1]), sigma_max=stats['sigma_max'],
                      note='per-rank slabs share one lattice plane; vertices on it are emitted by both neighbours (not index-welded across ranks)'),
            e2e=dict(value=med(t_e2e), unit='ms', h2d_bytes_per_step=17 * 4, d2h_bytes_per_step=d2h,
                     path='camera H2D -> rays -> render -> gather -> frame D2H; density grid -> marching cubes -> verts/faces/normals D2H'),
            gpu_launches=K * (2 + (per // 32768 + 1) * 7 + ((x1 - x0) // args.mesh_slab + 1) * 3 + 6), clocks=clock_info,
            roofline=dict(kernel='density pass over the lattice (arcb200_world_to_unit + grid encode + geometry head per slab)', bound='hbm',
                          achieved=algo / (grid_ms * 1e-3) / 1e9, peak=peaks['hbm_gbs'], unit='GB/s',
                          frac=algo / (grid_ms * 1e-3) / 1e9 / peaks['hbm_gbs'], traffic=None,
                          algorithmic_bytes_per_launch=algo,
                          note='gathers are served by L2 (the table is 25 MB); the fraction is against the HBM copy peak, '
                               'see profiles/l2_gather_peak.json for the L2 sector roof of the gather itself'),
            cpu_baseline=cpu)))


def _cpu_arm(rp, n_grid, level, n_rays_frame):
    """bounded sample on the host: the oracle's NGP forward on 512 rays, its field (grid encode + both heads: a slight over-count
    of the density-only pass) on 32^3 lattice points, the numpy marching cubes on the 128^3 grid of the same field; extrapolated
    linearly to the full frame / lattice"""
    from arcnerf_b200 import synth
    from oracle import marching_cubes as M
    from oracle.ngp import NGP_CONFIG, NGPOracle
    orc = NGPOracle(dict(table=rp.table.cpu().numpy(), geo=rp.geo_p.cpu().numpy(), rad=rp.rad_p.cpu().numpy()), NGP_CONFIG)
    orc.set_occupancy(rp.volume.cpu().numpy().astype(bool))
    o, d = synth.camera_rays(32, 16, synth.sphere_poses(1, seed=5)[0])
    t0 = time.time()
    orc.forward(o, d)
    t_rays = time.time() - t0
    lo, hi = rp.aabb[0].tolist(), rp.aabb[1].tolist()
    unit = ((M.volume_pts(lo, hi, 32) - np.asarray(lo, np.float32)) / (np.asarray(hi, np.float32) - np.asarray(lo, np.float32))).astype(np.float32)
    dirs = np.tile(np.array([[0.0, 0.0, 1.0]], np.float32), (unit.shape[0], 1))
    t0 = time.time()
    orc.field(unit, dirs)
    t_pts = time.time() - t0
    g = rp.density_grid(128).cpu().numpy()
    t0 = time.time()
    M.extract_mesh(g, level, [2.0 / 128] * 3, [2.0] * 3)
    t_mc = time.time() - t0
    full = t_rays * n_rays_frame / 512 + t_pts * n_grid ** 3 / 32 ** 3 + t_mc * (n_grid / 128) ** 3
    return dict(kind='port', cores=1, unit='ms', value=full * 1e3,
                sample='512 rays (%.2f s), field at 32^3 lattice points (%.2f s), numpy marching cubes on 128^3 (%.2f s); '
                       'extrapolated linearly to 2 073 600 rays and %d^3' % (t_rays, t_pts, t_mc, n_grid))
