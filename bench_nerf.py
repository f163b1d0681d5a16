"""`python bench.py --config nerf`: the vanilla-NeRF training step (BASELINE.json configs[0]) on the dense-chain kernels.

A step = 1024 rays: 64 stratified coarse samples through the coarse networks, compositing, 128 inverse-CDF samples, 192 merged
samples through the fine networks, compositing, MSE on both stages, backward through everything, Adam on both network pairs
(configs/models/nerf.yaml + configs/expr/NeRF/lego/nerf_lego_nerf.yaml).  262 144 network evaluations of 1.19 MFLOP each
forward, three times that with the backward: 0.94 TFLOP per step.
--impl reference / cpu_baseline: the same step through oracle/nerf.py (torch CPU fp32 restatement of the reference's NeRF
model, pinned to the reference's own FullModel by tests/golden/models_golden.npz) on the box's host cores.
"""
import json
import os
import time

import numpy as np
import torch

METRIC = 'nerf_train_rays_per_sec'
UNIT = 'rays/s'
N_RAYS = 1024
WORKLOAD = ('vanilla NeRF (BASELINE configs[0]): frequency embedding (10 / 4 bands), 8x256 geometry MLP with skip + 128-wide '
            'radiance head, 64 coarse + 128 fine samples per ray, 1024-ray batch, separate coarse and fine networks, '
            'Blender-shaped synthetic rays, random-init weights')
NERF_CFG = dict(
    rays=dict(near=2.0, far=6.0, n_sample=64, n_importance=128, inverse_linear=False, perturb=True, add_inf_z=True, noise_std=0.0,
              white_bkg=True),
    chunk_rays=32768,
    geometry=dict(W=256, D=8, skips=[4], encoder=dict(type='FreqEmbedder', input_dim=3, n_freqs=10), W_feat=256, geometric_init=False),
    radiance=dict(mode='vf', W=128, D=1, encoder=dict(view=dict(type='FreqEmbedder', input_dim=3, n_freqs=4)), W_feat_in=256))
FLOP_PER_POINT = 2.0 * (63 * 256 + 6 * 256 * 256 + 319 * 256 + 256 * 257 + 283 * 128 + 128 * 3)
POINTS_PER_RAY = 64 + 192
A100_S_PER_ITER = 0.25          # docs/expr.md:130 (1x A100)


def _batches(n, seed):
    from arcnerf_b200 import synth
    out = []
    for b in range(n):
        o, d = synth.train_ray_batch(N_RAYS, seed=seed * 1000 + b)
        tgt = np.random.RandomState(seed * 1000 + b + 7).uniform(0, 1, (N_RAYS, 3)).astype(np.float32)
        out.append(tuple(torch.from_numpy(a) for a in (o, d, tgt)))
    return out


def cpu_step_fn(threads):
    """one optimisation step of the oracle's NeRF (torch CPU): returns step(o, d, target)"""
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), 'tests'))
    import nerf_full_cases as nc
    from oracle import nerf as on
    torch.set_num_threads(threads)
    coarse = on.Nets(*nc.weights('nerf', 10), requires_grad=True)
    fine = on.Nets(*nc.weights('nerf', 11), requires_grad=True)
    opt = torch.optim.Adam(coarse.parameters() + fine.parameters(), lr=5e-4)
    cfg = dict(NERF_CFG['rays'], perturb=False)

    def step(o, d, tgt):
        out_c, out_f = on.nerf_forward(coarse, fine, o, d, cfg)
        loss = ((out_c['rgb'] - tgt) ** 2).mean() + ((out_f['rgb'] - tgt) ** 2).mean()
        opt.zero_grad()
        loss.backward()
        opt.step()
        return float(loss.detach())
    return step


def run_reference(args):
    threads = os.cpu_count() or 1
    step = cpu_step_fn(threads)
    rays = args.ref_rays if args.ref_rays and args.ref_rays < N_RAYS else 256      # a bounded sample: the CPU step is ~1 s per 100 rays
    b = _batches(args.steps + args.warmup, 0)
    for i in range(args.warmup):
        step(*(x[:rays] for x in b[i]))
    t0 = time.perf_counter()
    for i in range(args.steps):
        step(*(x[:rays] for x in b[args.warmup + i]))
    dt = time.perf_counter() - t0
    v = rays * args.steps / dt
    print(json.dumps(dict(impl='reference', metric=METRIC, value=v, unit=UNIT, n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
                          ms_per_step=1e3 * dt / args.steps, higher_is_better=True, scaling='weak', vs_baseline=None,
                          dtype='f32 (CPU)', data='synthetic', config=dict(workload=WORKLOAD, rays_per_step=rays),
                          cpu_baseline=dict(value=v, unit=UNIT, cores=threads, kind='port',
                                            sample='%d steps of %d rays (the GPU arm steps 1024) through oracle/nerf.py, the torch-CPU '
                                                   'restatement of the reference NeRF model' % (args.steps, rays)),
                          e2e=dict(value=v, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0))))


def run(args, rank, world, dev, dist, read_peaks, ClockSampler):
    from arcnerf_b200 import functional as F
    from arcnerf_b200.plugin.nerf_model import NeRF
    # multi-rank: NCCL sets up its channels per message size during the first collectives - keep them out of the timed region
    K, W = args.steps, max(args.warmup, 3 if world == 1 else 10)
    torch.manual_seed(1337 + rank)
    model = NeRF(NERF_CFG, device=str(dev))
    nets = [model.coarse_nets, model.fine_nets]
    moments = [(torch.zeros_like(n.flat), torch.zeros_like(n.flat)) for n in nets]
    host = [tuple(x.pin_memory() for x in b) for b in _batches(K + W, rank)]
    resident = [tuple(x.to(dev) for x in b) for b in host]
    step_no = [0]

    def train_step(o, d, tgt):
        out = model({'rays_o': o, 'rays_d': d})
        loss = ((out['rgb_coarse'] - tgt) ** 2).mean() + ((out['rgb_fine'] - tgt) ** 2).mean()
        for n in nets:
            n.flat.grad = None
        loss.backward()
        step_no[0] += 1
        for n, (m, v) in zip(nets, moments):
            g = n.flat.grad
            if world > 1:
                dist.all_reduce(g)
            F.adam_step(n.flat.data, m, v, g, step_no[0], 5e-4, grad_scale=1.0 / world, zero_grad=False)
            n.flat.data_ptr()
            n._packed_version = None            # parameters changed in place through the library: re-pack at the next forward
        return loss

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
    value = N_RAYS * world * K / (ms_total * 1e-3)
    # end to end from pinned host buffers: H2D of rays + targets, loss D2H + sync, every step
    from bench_common import LossReader
    reader = LossReader()
    dbuf = [torch.empty_like(x, device=dev) for x in host[0]]
    barrier()
    e0.record()
    for i in range(K):
        for dst, src in zip(dbuf, host[W + i]):
            dst.copy_(src, non_blocking=True)
        loss = train_step(*dbuf)
        reader.push(loss)
    reader.finish()
    e1.record()
    barrier()
    ms_e2e = max_over_ranks(e0.elapsed_time(e1))
    clock_info = clocks.stop()
    # the dominant kernels, timed alone with CUDA events: the fine pass of the dense chain forward (+ save) and backward
    from arcnerf_b200 import _lib
    fine = model.fine_nets
    n = N_RAYS * 192
    x = torch.rand((n, 3), device=dev) * 2 - 1
    dd = torch.randn((n, 3), device=dev)
    gs, gc = torch.randn((n, 1), device=dev), torch.randn((n, 3), device=dev)

    def fb():
        s, c = fine(x, dd)
        torch.autograd.backward([s, c], [gs, gc])
        fine.flat.grad = None
    for _ in range(3):
        fb()
    torch.cuda.synchronize()
    ts = []
    for _ in range(10):
        e0.record(); fb(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms_fb = float(np.median(ts))
    peaks = read_peaks()
    flops_fb = 3.0 * FLOP_PER_POINT * n
    ach = flops_fb / (ms_fb * 1e-3) / 1e12
    roofline = dict(kernel='dense_chain_kernel<fwd> + dense_chain_kernel<bwd> + dense_wgrad_kernel (fine pass: 196 608 points)', bound='tensor',
                    achieved=ach, peak=peaks['bf16_tflops'], unit='TFLOP/s', frac=ach / peaks['bf16_tflops'], traffic=None,
                    flops_per_launch_group=flops_fb, ms_per_launch_group=ms_fb, peak_source=peaks['source'],
                    whole_step_tflops=3.0 * FLOP_PER_POINT * POINTS_PER_RAY * N_RAYS / (ms_total / K * 1e-3) / 1e12)
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        step = cpu_step_fn(threads)
        rays = 128
        b = _batches(6, 99)
        step(*(t[:rays] for t in b[0]))
        t0 = time.perf_counter()
        n_cpu = 0
        while n_cpu < 5 and time.perf_counter() - t0 < 20.0:
            step(*(t[:rays] for t in b[1 + n_cpu]))
            n_cpu += 1
        dt = time.perf_counter() - t0
        cpu_baseline = dict(value=rays * n_cpu / dt, unit=UNIT, cores=threads, kind='port',
                            sample='%d steps of %d rays (the GPU arm steps 1024) through oracle/nerf.py: torch-CPU fp32 NeRF '
                                   '(coarse + fine networks, compositing, inverse-CDF re-sampling, MSE, autograd, Adam)' % (n_cpu, rays))
    if rank == 0:
        h2d = sum(t.numel() * t.element_size() for t in host[0])
        print(json.dumps(dict(
            metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=K, warmup=W, ms_per_step=ms_total / K, higher_is_better=True,
            scaling='weak', vs_baseline=value / (N_RAYS / A100_S_PER_ITER) if world == 1 else None,
            baseline_note='reference publishes 0.25 s/iter for vanilla NeRF on 1x A100 (docs/expr.md:130) = 4096 rays/s at 1024 rays',
            dtype='fp16 operands / fp32 accumulate', data='synthetic',
            config=dict(workload=WORKLOAD, rays_per_step_per_gpu=N_RAYS, points_per_step_per_gpu=N_RAYS * POINTS_PER_RAY,
                        l2='every step streams 2.5 GB of saved activations and gradients (fp16) through HBM, far beyond the 126 MB L2',
                        parallelism='dp%d' % world, gradient_exchange='NCCL all-reduce of the two flat parameter gradients' if world > 1 else 'single GPU'),
            e2e=dict(value=N_RAYS * world * K / (ms_e2e * 1e-3), unit=UNIT, h2d_bytes_per_step=h2d, d2h_bytes_per_step=4, ms_per_step=ms_e2e / K,
                     path='host pinned buffers -> H2D per step -> NeRF.forward + backward + Adam -> loss D2H every step, read by the host one step behind'),
            gpu_launches=K * (2 * (1 + 1 + 1 + 1 + 1) + 6 + 4 + 2 * 3),      # per step: 2 x (chain fwd, chain bwd, wgrad, pack x2..), zvals/points/composite/sample kernels, Adam
            clocks=clock_info, roofline=roofline, cpu_baseline=cpu_baseline)))
