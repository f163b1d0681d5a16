"""`python bench.py --config neus`: the NeuS-NGP training step (BASELINE.json configs[2]; configs/expr/NeRF/lego/nerf_lego_neus_ngp.yaml)
data-parallel over the ranks: 4096 rays per GPU through the occupancy march, hash grid with input gradients, the 64-wide SDF
network, the 'pvnf' radiance head, sdf_to_alpha and compositing; Huber colour loss x5 + eikonal loss x0.1 (second-order backward
through the grid and the SDF network); gradients averaged over ranks with NCCL (what DDP does in the reference,
common/trainer/basic_trainer.py:192-202); Adam on every parameter tensor with the library's kernel.
No CPU arm: the reference cannot run this configuration without a GPU (tinycudann), and the composed oracle (oracle/neus.py)
is forward-only; the line carries cpu_baseline = null and says so."""
import json
import os

import numpy as np
import torch

METRIC = 'neus_ngp_train_rays_per_sec'
UNIT = 'rays/s'
N_RAYS = 4096
WORKLOAD = ('NeuS-NGP (BASELINE configs[2]): 128^3 occupancy march, 16-level hash grid with dy/dx, 32-64-17 weight-normed softplus SDF '
            'network with normals, pvnf fused radiance head, sdf_to_alpha + compositing, Huber + eikonal loss (double backward), '
            '4096-ray batch per GPU, lego-like 7% occupancy, data-parallel with NCCL gradient all-reduce')
A100_S_PER_ITER = 0.06          # docs/expr.md:150 (1x A100)


def run(args, rank, world, dev, dist, read_peaks, ClockSampler):
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), 'tests'))
    from arcnerf_b200 import functional as F
    from arcnerf_b200 import synth
    from arcnerf_b200.plugin.neus_model import Neus, eikonal_loss
    from test_neus_gpu import CFG
    # multi-rank: NCCL sets up its channels per message size during the first collectives - keep them out of the timed region
    K, W = args.steps, max(args.warmup, 3 if world == 1 else 10)
    torch.manual_seed(1337)                                  # identical replicas on every rank
    model = Neus(CFG, device=str(dev))
    model.obj_bound.volume.bitfield.copy_(torch.from_numpy(synth.occupancy_volume('lego')).to(dev))
    params = [p for p in model.parameters() if p.requires_grad]
    moments = [(torch.zeros_like(p), torch.zeros_like(p)) for p in params]
    batches = []
    for b in range(K + W):
        o, d = synth.train_ray_batch(N_RAYS, seed=rank * 1000 + b)
        rs = np.random.RandomState(rank * 1000 + b + 7)
        batches.append(tuple(torch.from_numpy(a).pin_memory() for a in (o, d, rs.uniform(0, 1, (N_RAYS, 3)).astype(np.float32),
                                                                         rs.uniform(0, 1, (N_RAYS, 3)).astype(np.float32))))
    resident = [tuple(x.to(dev) for x in b) for b in batches]
    step_no = [0]

    def train_step(o, d, tgt, bkg):
        out = model({'rays_o': o, 'rays_d': d, 'bkg_color': bkg}, cur_epoch=step_no[0])
        loss = torch.nn.functional.huber_loss(out['rgb'], tgt, delta=0.1) * 5.0 + 0.1 * eikonal_loss(out['normal_pts'])
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
            F.adam_step(p.data.view(-1), m.view(-1), v.view(-1), g.view(-1), step_no[0], 1e-3, eps=1e-15, grad_scale=1.0 / world, zero_grad=False)
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
        h2d = sum(t.numel() * t.element_size() for t in batches[0])
        n_table = model.geo_net.embed_fn.hashgrid_encode_tcnn.params.numel()
        peaks = read_peaks()
        print(json.dumps(dict(
            metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=K, warmup=W, ms_per_step=ms_total / K, higher_is_better=True,
            scaling='weak', vs_baseline=value / (N_RAYS / A100_S_PER_ITER) if world == 1 else None,
            baseline_note='reference publishes 0.06 s/iter for NeuS-NGP on 1x A100 (docs/expr.md:150) = 68 267 rays/s',
            dtype='fp16 table and radiance head, fp32 SDF network and gradients', data='synthetic',
            config=dict(workload=WORKLOAD, rays_per_step_per_gpu=N_RAYS, parallelism='dp%d' % world,
                        l2='per-step working set (hash-table gradient + Adam state, 0.2 GB) exceeds the 126 MB L2; every step draws new rays',
                        gradient_exchange='NCCL all-reduce per parameter tensor (48.8 MB hash-table gradient + 5 small tensors)' if world > 1 else 'single GPU'),
            e2e=dict(value=N_RAYS * world * K / (ms_e2e * 1e-3), unit=UNIT, h2d_bytes_per_step=h2d, d2h_bytes_per_step=4, ms_per_step=ms_e2e / K,
                     path='host pinned buffers -> H2D per step -> Neus.forward + losses + backward (incl. second order) + Adam -> loss D2H every step, read by the host one step behind'),
            gpu_launches=K * 60,
            clocks=clock_info,
            roofline=dict(kernel='adam_kernel on the hash table (dense sweep, the largest fixed cost of the step)', bound='hbm', achieved=None,
                          peak=peaks['hbm_gbs'], unit='GB/s', frac=None, traffic=None, algorithmic_bytes_per_launch=32.0 * n_table,
                          note='this configuration is a composition of the library kernels under torch autograd (many small launches); '
                               'per-kernel rooflines are reported by the ngp and nerf configurations'),
            cpu_baseline=None, cpu_baseline_note='the reference cannot run NeuS-NGP without a GPU (tinycudann); no CPU arm')))
