#!/usr/bin/env python
"""Benchmark of the NGP hot path (BASELINE.json configs[1]): NGP training rays/s (+ 800x800 frame render ms).

    python bench.py [--gpus N --steps K --warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

A step = one optimisation step of the NGP foreground model on one 4096-ray batch per GPU:
slab test + occupancy march + hash-grid encode + fused MLPs + compositing + Huber loss + backward + (NCCL all-reduce
of the gradients when N > 1) + Adam.  Rank 0 prints ONE JSON line.
  value          : rays/s with the ray batches already resident in HBM (whole job, all ranks)
  e2e            : the same through the public host API with HOST (pinned) ray buffers: H2D of rays/targets and D2H of
                   the loss inside the timed region, every step
  roofline       : dominant kernel of the step, timed live with CUDA events on the launching stream
  cpu_baseline   : the CPU oracle port of the same step (oracle/ngp.py) timed on this box's host cores (rank 0, N=1)
  render_ms      : 800x800 frame (640 000 rays), forward only, rays sharded over ranks + final gather
--impl reference : times the CPU implementation of the same step (the oracle port; the reference itself cannot run
                   this config on CPU) and prints the same line with "impl": "reference".
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

METRIC = 'ngp_train_rays_per_sec'
UNIT = 'rays/s'
N_RAYS = 4096
WORKLOAD = ('NGP (BASELINE configs[1]): 16-level hash grid (2^19 x 2 feat) + 64-wide fused MLPs + 128^3 occupancy '
            'pruning, 4096-ray batch per GPU, Blender-shaped 800x800 synthetic rays, lego-like 7% occupancy, '
            'random-init weights')
BASELINE_A100_S_PER_ITER = 0.018   # docs/expr.md:136 (1x A100, 4096-ray initial batch)


def read_peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm_gbs=float(d['hbm_gbs']), bf16_tflops=float(d.get('bf16_tflops_sustained', d['bf16_tflops'])),
                    source='measured (MEASURED_PEAKS.json)')
    return dict(hbm_gbs=6650.0, bf16_tflops=1400.0, source='fallback (B200_PROFILING.md)')


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region.  index None: no sampling (ranks other
    than 0 - eight nvidia-smi pollers beside eight host-bound training loops cost more than they tell)."""

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        if self.index is None:
            return
        q = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
             'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + q,
                                          '--format=csv,noheader,nounits', '-lms', '200'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=['nvidia-smi unavailable'])
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(',')]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap'), f[2:6]):
                if v.lower().startswith('active'):
                    reasons.add(name)
        return dict(sm_mhz=statistics.median(sm) if sm else None, sm_max_mhz=max(mx) if mx else None,
                    reasons=sorted(reasons), samples=len(sm))


def make_batches(n_batches, seed):
    """host (pinned) ray batches: rays_o, rays_d, target rgb, random background colour (the reference trains with a
    random bkg colour: configs/expr/NeRF/lego/nerf_lego_nerf_ngp.yaml dataset.train.scheduler.bkg_color)"""
    from arcnerf_b200 import synth
    out = []
    for b in range(n_batches):
        o, d = synth.train_ray_batch(N_RAYS, seed=seed * 1000 + b)
        rng = np.random.RandomState(seed * 1000 + b + 7)
        tgt = rng.uniform(0, 1, (N_RAYS, 3)).astype(np.float32)
        bkg = rng.uniform(0, 1, (N_RAYS, 3)).astype(np.float32)
        out.append(tuple(torch.from_numpy(a).pin_memory() if torch.cuda.is_available() else torch.from_numpy(a)
                         for a in (o, d, tgt, bkg)))
    return out


EXCHANGE_NOTE = {
    'none': 'single GPU',
    'peer': 'fused into the Adam kernel: peer loads of the gradient slices + peer stores of the fp16 table over NVLink (csrc/peer.cu)',
    'peer16': 'fused into the Adam kernel over NVLink peer memory (csrc/peer.cu): the hash-table gradient travels as a loss-scaled fp16 copy (the precision tcnn hands to DDP in the reference), summed in fp32 with the owner\'s own fp32 slice; fp16 table slices stored back to every rank',
    'mc16': 'done by the NVSwitch, fused into the Adam kernel (csrc/peer.cu adam_mc16_kernel): one multimem.ld_reduce per 8 entries sums every rank\'s loss-scaled fp16 gradient copy in the fabric (fp32 accumulation), one multimem.st broadcasts the updated fp16 table entries to every rank',
    'nccl': 'one NCCL all-reduce of the flat fp32 gradient buffer per step, then Adam on every rank',
}


# algorithmic work per unit (SURVEY.md §8d / DESIGN.md §4)
def stage_work(name, n_samples, n_rays, n_table_params, world=1):
    if name == 'grid_fwd':
        return 'hbm', 588.0 * n_samples          # 12 B in + 16 levels * 8 corners * 4 B gathered + 64 B out
    if name == 'field_fwd':                      # the gathers bound it (L1/L2 sector rate); MLP flops 20480/sample ride along
        return 'hbm', (588.0 + 16.0) * n_samples  # encode bytes + sigma, rgb out
    if name == 'grid_bwd':
        return 'hbm', 588.0 * n_samples          # 12 B + 64 B in + 16*8 entries of 4 B updated
    if name == 'adam_table':
        return 'hbm', 34.0 * n_table_params      # p,m,v,g read; p,m,v,g(=0) written; fp16 copy written
    if name == 'adam_table_peer':               # per rank: its 1/world slice; p,m,v in and out, `world` gradients in, `world` fp16 copies out
        return 'hbm', (24.0 + 6.0 * world) * n_table_params / world
    if name == 'composite_train':                # sigma, rgb, t read (twice, second time from L1/L2); d_sigma, d_rgb written
        return 'hbm', 36.0 * n_samples + 44.0 * n_rays
    if name == 'composite_fwd':
        return 'hbm', 20.0 * n_samples + 20.0 * n_rays
    if name == 'composite_bwd':
        return 'hbm', 36.0 * n_samples + 20.0 * n_rays
    if name == 'march_packed':                   # latency-bound by design (a warp walks its ray); bytes given for reference only
        return 'hbm', 24.0 * n_rays + 8.0 * n_samples
    if name in ('mlp_geo_fwd',):
        return 'tensor', 6144.0 * n_samples
    if name in ('mlp_rad_fwd',):
        return 'tensor', 14336.0 * n_samples
    if name == 'mlp_geo_bwd':
        return 'tensor', 2 * 6144.0 * n_samples   # dgrad + wgrad (hidden activations come saved from the forward)
    if name == 'mlp_rad_bwd':
        return 'tensor', 2 * 14336.0 * n_samples
    return None, 0.0


def run_reference_arm(args, rank, world):
    """CPU implementation of the same step (oracle port) on the host cores; rank 0 only."""
    if rank != 0:
        return
    from arcnerf_b200 import synth                       # numpy-only input generators; the product library is not loaded
    from oracle.ngp import NGP_CONFIG, NGPOracle, init_params
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    cfg = dict(NGP_CONFIG)
    params = init_params(cfg, seed=1337)
    orc = NGPOracle({k: v.numpy() for k, v in params.items()}, cfg, threads=threads)
    orc.set_occupancy(synth.occupancy_volume('lego'))
    sample_rays = args.ref_rays
    batches = make_batches(args.steps + args.warmup, seed=0)
    for i in range(args.warmup):
        o, d, t, b = (x.numpy()[:sample_rays] for x in batches[i])
        orc.train_step(o, d, t, b)
    t0 = time.perf_counter()
    for i in range(args.steps):
        o, d, t, b = (x.numpy()[:sample_rays] for x in batches[args.warmup + i])
        orc.train_step(o, d, t, b)
    dt = time.perf_counter() - t0
    value = sample_rays * args.steps / dt
    line = dict(impl='reference', metric=METRIC, value=value, unit=UNIT, n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
                ms_per_step=1e3 * dt / args.steps, higher_is_better=True, scaling='weak', vs_baseline=None,
                dtype='fp16 storage / fp32 accumulate (CPU emulation)', data='synthetic',
                config=dict(workload=WORKLOAD, rays_per_step=sample_rays),
                cpu_baseline=dict(value=value, unit=UNIT, cores=threads, kind='port',
                                  sample='%d steps of %d rays (the GPU arm steps 4096 rays) through oracle/ngp.py: '
                                         'C sampler + torch-CPU hash grid / MLPs / ray_marching + torch Adam' % (args.steps, sample_rays)),
                e2e=dict(value=value, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    print(json.dumps(line))


def main():
    if os.environ.get('ARCB200_BENCH_WATCHDOG'):        # debugging aid: dump every thread's stack and exit if the run hangs
        import faulthandler
        faulthandler.dump_traceback_later(float(os.environ['ARCB200_BENCH_WATCHDOG']), exit=True)
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=50)
    ap.add_argument('--warmup', type=int, default=10)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--ref-rays', type=int, default=N_RAYS, help='rays per step of the CPU arm (default: the same 4096-ray step as the GPU arm)')
    ap.add_argument('--no-render', action='store_true')
    ap.add_argument('--occ-every', type=int, default=16, help='occupancy refresh period in steps (0 = never); the reference config uses 16')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-prefetch', action='store_true', help='march every batch on the main stream (measures what the side-stream sampler prefetch hides)')
    ap.add_argument('--stages', action='store_true', help='print the per-stage timing table to stderr')
    ap.add_argument('--config', default='ngp', choices=['ngp', 'nerf', 'neus', 'ngp_nerfpp', 'mesh'],
                    help="ngp: BASELINE configs[1] (the headline); nerf: the vanilla-NeRF step of configs[0] (bench_nerf.py); neus: the NeuS-NGP step "
                         "of configs[2] (bench_neus.py); ngp_nerfpp: NGP foreground + NeRF++ background training of configs[3]; mesh: 1920x1080 "
                         "frame + 512^3 mesh extraction of configs[4] (both bench_scene.py)")
    ap.add_argument('--mesh-grid', type=int, default=512)
    ap.add_argument('--mesh-slab', type=int, default=16)
    args = ap.parse_args()

    rank = int(os.environ.get('RANK', 0))
    local_rank = int(os.environ.get('LOCAL_RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))

    if args.impl == 'reference':
        if args.config == 'nerf':
            if rank == 0:
                import bench_nerf
                bench_nerf.run_reference(args)
            return
        if args.config in ('ngp_nerfpp', 'mesh'):
            if rank == 0:
                print(json.dumps(dict(impl='reference', unavailable='the reference needs tinycudann and skimage (GPU / absent packages) for this configuration; '
                                                                    'the CPU port of its parts is timed inside the ours line (cpu_baseline)')))
            return
        if args.config == 'neus':
            if rank == 0:
                print(json.dumps(dict(impl='reference', unavailable='NeuS-NGP needs tinycudann on a GPU in the reference; no CPU implementation of this configuration exists')))
            return
        run_reference_arm(args, rank, world)
        return

    if not torch.cuda.is_available():
        raise RuntimeError('bench.py needs a CUDA device: the hot path has no CPU fallback')
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group('nccl', device_id=torch.device('cuda', local_rank))
    else:
        dist = None
        torch.cuda.set_device(0)
    dev = torch.device('cuda', local_rank if world > 1 else 0)

    if args.config in ('nerf', 'neus', 'ngp_nerfpp', 'mesh'):
        if args.config in ('ngp_nerfpp', 'mesh'):
            import bench_scene
            getattr(bench_scene, 'run_' + args.config)(args, rank, world, dev, dist, read_peaks, ClockSampler)
        else:
            mod = __import__('bench_' + args.config)
            mod.run(args, rank, world, dev, dist, read_peaks, ClockSampler)
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return

    from arcnerf_b200 import synth
    from arcnerf_b200.ngp import NGPPipeline, StageTimer

    K, W = args.steps, max(args.warmup, 3)
    pipe = NGPPipeline(max_rays=N_RAYS, capacity=1 << 19, device=str(dev), seed=1337, world_size=world)
    pipe.set_occupancy(synth.occupancy_volume('lego'))
    host = make_batches(K + W, seed=rank)
    resident = [tuple(x.to(dev) for x in b) for b in host]
    torch.cuda.synchronize()

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

    # ---- warm-up
    for i in range(W):
        o, d, t, b = resident[i]
        pipe.train_step(o, d, t, b)
    barrier()

    clocks = ClockSampler(dev.index if rank == 0 else None)
    clocks.start()
    # ---- timed: resident inputs
    occ_every = args.occ_every
    occ_calls = [0]

    def maybe_refresh(step_index):
        """VolumeBound.optimize every `occ_every` steps in its steady-state form (n/4 random + n/4 occupied voxels of 128^3,
        density through the field kernels).  All of the work runs inside the timed region; the refreshed occupancy is NOT
        handed to the sampler, because a random-init field would replace the lego-like synthetic scene by noise and change
        the sample count per ray."""
        if occ_every > 0 and (step_index + 1) % occ_every == 0:
            occ_calls[0] += 1
            pipe.update_occupancy(256 + 16 * occ_calls[0], apply=False)

    if occ_every > 0:
        pipe.update_occupancy(256, apply=False)      # untimed: allocates the refresh buffers
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for i in range(K):
        o, d, t, b = resident[W + i]
        nxt = resident[W + i + 1][:2] if (i + 1 < K and not args.no_prefetch) else None      # the trainer knows its next batch: sample it meanwhile
        pipe.train_step(o, d, t, b, next_rays=nxt)
        maybe_refresh(i)
    e1.record()
    barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    n_samples = int(pipe.meta[0])
    value = N_RAYS * world * K / (ms_total * 1e-3)

    # ---- timed: end to end from host buffers (H2D rays + targets, D2H loss, every step)
    loss_host = torch.empty(1, dtype=torch.float32).pin_memory()
    use_graph = os.environ.get('ARCB200_BENCH_GRAPH', '1') == '1'
    if use_graph:
        # static input slots + CUDA-graph replay of the post-sampler part of the step (incl. the NCCL all-reduce)
        pipe.enable_graphs(N_RAYS)
        for i in range(3):
            pipe.train_step_graphed(host[i], next_batch=host[i + 1])
        barrier()
        e0.record()
        # the caller reads every step's loss, one step behind: the D2H copy of step i is queued behind step i, and the host waits
        # for it only after step i + 1 has been queued (a training loop that logs its loss does not need to stall the device for it)
        loss_slots = [loss_host, torch.empty(1, dtype=torch.float32).pin_memory()]
        loss_ready = [torch.cuda.Event(), torch.cuda.Event()]
        losses_read = 0
        for i in range(K):
            nb = host[W + i + 1] if i + 1 < K else None
            loss = pipe.train_step_graphed(host[W + i], next_batch=nb)     # H2D of rays/targets inside
            loss_slots[i % 2].copy_(loss, non_blocking=True)
            loss_ready[i % 2].record()
            maybe_refresh(i)
            if i > 0:
                loss_ready[(i - 1) % 2].synchronize()
                losses_read += int(np.isfinite(float(loss_slots[(i - 1) % 2][0])))
        loss_ready[(K - 1) % 2].synchronize()
        losses_read += int(np.isfinite(float(loss_slots[(K - 1) % 2][0])))
        assert losses_read == K, 'a step produced a non-finite loss'
        e1.record()
        barrier()
        e2e_mode = ('host pinned buffers -> static device slots (H2D per step) -> sampler on side stream + CUDA graph replay -> loss D2H '
                    'every step, read by the host one step behind')
    else:
        dbufs = [[torch.empty_like(x, device=dev) for x in host[0]] for _ in range(2)]   # current / next batch
        barrier()
        e0.record()
        for dst, src in zip(dbufs[0], host[W]):
            dst.copy_(src, non_blocking=True)
        for i in range(K):
            cur, nxt = dbufs[i % 2], dbufs[(i + 1) % 2]
            if i + 1 < K:                                   # upload the next batch, then let the sampler run ahead on it
                for dst, src in zip(nxt, host[W + i + 1]):
                    dst.copy_(src, non_blocking=True)
            loss = pipe.train_step(*cur, next_rays=(nxt[0], nxt[1]) if i + 1 < K else None)
            loss_host.copy_(loss, non_blocking=True)
            maybe_refresh(i)
            torch.cuda.current_stream().synchronize()      # the caller reads the loss every step
        e1.record()
        barrier()
        e2e_mode = 'host pinned buffers -> H2D per step -> eager launches (gradient all-reduce is not graph captured) -> loss D2H + sync per step'
    ms_e2e = max_over_ranks(e0.elapsed_time(e1))
    e2e_value = N_RAYS * world * K / (ms_e2e * 1e-3)
    h2d = sum(x.numel() * x.element_size() for x in host[0])
    clock_info = clocks.stop()

    # ---- per-stage CUDA-event pass (same stream, same work) for the roofline of the dominant kernel
    timer = StageTimer()
    pipe.timer = timer
    for i in range(min(K, 20)):
        o, d, t, b = resident[W + i]
        pipe.train_step(o, d, t, b)
    torch.cuda.synchronize()
    pipe.timer = None
    stages = timer.summary()
    peaks = read_peaks()
    n_table = pipe.table.numel()
    top = max(stages.items(), key=lambda kv: kv[1][0])
    step_ms_staged = sum(v[0] for v in stages.values())
    table = []
    for name, (ms, cnt) in sorted(stages.items(), key=lambda kv: -kv[1][0]):
        bound, work = stage_work(name, n_samples, N_RAYS, n_table, world)
        ach = work / (ms * 1e-3) / (1e9 if bound == 'hbm' else 1e12) if bound else None
        peak = peaks['hbm_gbs'] if bound == 'hbm' else (peaks['bf16_tflops'] if bound == 'tensor' else None)
        table.append(dict(stage=name, ms=round(ms, 4), share=round(ms / step_ms_staged, 3), bound=bound,
                          achieved=None if ach is None else round(ach, 2), frac=None if ach is None else round(ach / peak, 4)))
    if args.stages and rank == 0:
        for r in table:
            sys.stderr.write('%-16s %8.4f ms  share %5.3f  %-6s achieved %-10s frac %s\n' %
                             (r['stage'], r['ms'], r['share'], r['bound'], r['achieved'], r['frac']))
    bound, work = stage_work(top[0], n_samples, N_RAYS, n_table, world)
    if bound is None:
        bound, work = 'hbm', 0.0
    ach = work / (top[1][0] * 1e-3) / (1e9 if bound == 'hbm' else 1e12)
    peak = peaks['hbm_gbs'] if bound == 'hbm' else peaks['bf16_tflops']
    traffic, traffic_src = None, None
    try:                                        # DRAM bytes per launch of that kernel from the committed ncu --set full capture
        with open(os.path.join(ROOT, 'profiles', 'r02_traffic.json')) as f:
            tj = json.load(f)
        if top[0] in tj:
            traffic, traffic_src = tj[top[0]]['traffic_bytes'], 'profiles/r02_traffic.json (ncu --set full, dram read + write per launch)'
    except (OSError, ValueError):
        pass
    roofline = dict(kernel=top[0], bound=bound, achieved=ach, peak=peak, unit='GB/s' if bound == 'hbm' else 'TFLOP/s',
                    frac=ach / peak, traffic=traffic, traffic_source=traffic_src, algorithmic_bytes_per_launch=work if bound == 'hbm' else None, ms_per_launch=top[1][0], share_of_step=top[1][0] / step_ms_staged,
                    peak_source=peaks['source'], stages=table)

    # ---- 800x800 frame render (forward only), rays sharded over ranks, final gather
    render_ms = None
    if not args.no_render:
        del resident
        rp = NGPPipeline(max_rays=32768, capacity=3 << 19, device=str(dev),
                         params=dict(table=pipe.gather_master_table(), geo=pipe.geo_p, rad=pipe.rad_p))
        rp.set_occupancy(synth.occupancy_volume('lego'))
        o, d = synth.camera_rays(800, 800, synth.sphere_poses(1, seed=5)[0])
        n_all = o.shape[0]
        # ray tiles of 256 are dealt round-robin to the ranks: a contiguous split gives the ranks that see the object many more
        # samples than the ranks that see the border, and the frame time is the slowest rank's
        from arcnerf_b200.dist import deal_tiles, gather_dealt
        idx, per = deal_tiles(n_all, rank, world)
        o_d, d_d = torch.from_numpy(o[idx.numpy()]).to(dev), torch.from_numpy(d[idx.numpy()]).to(dev)
        white = torch.ones((1, 3), device=dev)
        # the frame is assembled on every rank by ONE all-gather + one strided copy that un-deals the tiles (dist.gather_dealt)
        frame = torch.empty((per * world, 3), device=dev) if dist is not None else None
        for _ in range(2):
            rp.render(o_d, d_d, white)
        times = []
        for _ in range(5):
            barrier()
            e0.record()
            rgb, _, _ = rp.render(o_d, d_d, white)
            if dist is not None:
                gather_dealt(rgb, n_all, world, out=frame)
            e1.record()
            barrier()
            times.append(max_over_ranks(e0.elapsed_time(e1)))
        render_ms = statistics.median(times)

    # ---- CPU baseline (rank 0, N = 1 only): bounded sample of the same step through the oracle port
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle.ngp import NGP_CONFIG, NGPOracle, init_params
        threads = os.cpu_count() or 1
        torch.set_num_threads(threads)
        cfg = dict(NGP_CONFIG)
        orc = NGPOracle({k: v.numpy() for k, v in init_params(cfg, seed=1337).items()}, cfg, threads=threads)
        orc.set_occupancy(synth.occupancy_volume('lego'))
        o, d, t, b = (x.numpy()[:args.ref_rays] for x in host[0])
        orc.train_step(o, d, t, b)                        # warm-up (page faults, Adam state)
        # about 12 s of CPU work: time one step, then run as many more as fit (bounded between 4 and 60 steps)
        t0 = time.perf_counter()
        o, d, t, b = (x.numpy()[:args.ref_rays] for x in host[1])
        orc.train_step(o, d, t, b)
        one = time.perf_counter() - t0
        n_cpu = int(min(60, max(3, 12.0 / max(one, 1e-3))))
        for i in range(1, n_cpu):
            o, d, t, b = (x.numpy()[:args.ref_rays] for x in host[(1 + i) % len(host)])
            orc.train_step(o, d, t, b)
        dt = time.perf_counter() - t0
        cpu_baseline = dict(value=args.ref_rays * n_cpu / dt, unit=UNIT, cores=threads, kind='port',
                            sample='%d steps of %d rays (the GPU arm steps 4096 rays), oracle/ngp.py: C sampler + '
                                   'torch-CPU hash grid / MLPs / ray_marching + torch Adam over the 12.2 M-parameter table'
                                   % (n_cpu, args.ref_rays))

    if rank == 0:
        line = dict(
            metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=K, warmup=W, ms_per_step=ms_total / K,
            higher_is_better=True, scaling='weak',
            vs_baseline=value / (N_RAYS / BASELINE_A100_S_PER_ITER) if world == 1 else None,
            baseline_note='reference publishes 0.018 s/iter for this config on 1x A100 (docs/expr.md:136) = 227 556 rays/s',
            dtype='fp16 storage / fp32 accumulate', data='synthetic',
            config=dict(workload=WORKLOAD, rays_per_step_per_gpu=N_RAYS, samples_per_step_per_gpu=n_samples,
                        l2='per-step working set (fp32 table + Adam moments + gradients + fp16 copy, 0.27 GB) exceeds the '
                           '126 MB L2 and every step draws a different ray batch; the Adam sweep evicts L2 between steps',
                        occupancy_refresh=('every %d steps inside the timed region: 524288 + 524288 voxels of 128^3 re-evaluated (steady-state VolumeBound.optimize), result not applied to the synthetic scene' % occ_every) if occ_every > 0 else 'off',
                        parallelism='dp%d' % world, gradient_exchange=EXCHANGE_NOTE[pipe.exchange]),
            e2e=dict(value=e2e_value, unit=UNIT, h2d_bytes_per_step=h2d, d2h_bytes_per_step=4, ms_per_step=ms_e2e / K, path=e2e_mode),
            gpu_launches=(NGPPipeline.LAUNCHES_TRAIN + {'peer': 2, 'peer16': 3, 'mc16': 3}.get(pipe.exchange, 0)) * K
            + (NGPPipeline.LAUNCHES_REFRESH * (K // occ_every) if occ_every > 0 else 0),
            clocks=clock_info, roofline=roofline, cpu_baseline=cpu_baseline,
            render_ms_800x800=render_ms,
        )
        print(json.dumps(line))
    if dist is not None:
        # orderly teardown: captured graphs and the peer arena are released (collectively) before the communicator
        sys.stdout.flush()
        if not args.no_render:
            rp.close()
        pipe.close()
        dist.barrier()
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
