"""Data-parallel step over peer memory (csrc/peer.cu): the fused exchange+Adam kernel against the plain
sum-then-Adam path, and (on a box with >= 2 GPUs) two ranks stepping the NGP pipeline with exchange='peer' against
exchange='nccl' (= DDP's all-reduce followed by Adam, common/trainer/basic_trainer.py:192-202)."""
import ctypes as C
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ptrs(tensors):
    return (C.c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])


@pytest.mark.parametrize('world', [1, 2, 3, 4, 8])
@pytest.mark.parametrize('n,lo,hi', [(4096, 0, 4096), (10007, 1024, 10007), (1 << 20, 4096 * 3, 4096 * 77 + 2)])
def test_adam_step_peer_matches_sum_then_adam(world, n, lo, hi):
    """all 'ranks' are buffers on one device: the kernel only sees pointers"""
    from arcnerf_b200 import functional as F
    from arcnerf_b200.peer import adam_step_peer
    g = torch.Generator(device='cuda').manual_seed(world * 131 + n)
    grads = [torch.randn(n, device='cuda', generator=g) for _ in range(world)]
    p0 = torch.randn(n, device='cuda', generator=g)
    m0 = torch.randn(n, device='cuda', generator=g) * 0.1
    v0 = torch.rand(n, device='cuda', generator=g) * 0.01
    halves = [torch.full((n,), 7.0, dtype=torch.float16, device='cuda') for _ in range(world)]
    # reference: rank-ordered sum, then the single-GPU kernel on the range
    gsum = grads[0].clone()
    for w in range(1, world):
        gsum += grads[w]
    p_ref, m_ref, v_ref = p0.clone(), m0.clone(), v0.clone()
    h_ref = torch.full((n,), 7.0, dtype=torch.float16, device='cuda')
    F.adam_step(p_ref[lo:hi], m_ref[lo:hi], v_ref[lo:hi], gsum[lo:hi], 3, 1e-2, weight_decay=1e-3, p_half=h_ref[lo:hi],
                grad_scale=1.0 / world, zero_grad=False)
    p, m, v = p0.clone(), m0.clone(), v0.clone()
    before = [x.clone() for x in grads]
    adam_step_peer(_ptrs(grads), _ptrs(halves), world, lo, hi, p, m, v, 3, 1e-2, weight_decay=1e-3, grad_scale=1.0 / world)
    torch.cuda.synchronize()
    assert torch.equal(p, p_ref) and torch.equal(m, m_ref) and torch.equal(v, v_ref)
    for h in halves:
        assert torch.equal(h, h_ref)            # the slice in every copy, nothing outside the range
    for a, b in zip(grads, before):
        assert torch.equal(a, b)                # gradients are left for the peers


@pytest.mark.parametrize('world', [1, 2, 8])
def test_mlp_adam_pack_matches_separate_kernels(world):
    """one launch for both heads == (sum of the gradient sources) -> adam_step -> mlp_pack per head, bit for bit"""
    from arcnerf_b200 import functional as F
    specs = [F.MlpSpec(32, 64, 1, 16, 'relu', 'none'), F.MlpSpec(32, 64, 2, 3, 'relu', 'sigmoid')]
    g = torch.Generator(device='cuda').manual_seed(world)
    ps = [torch.randn(sp.n_params, device='cuda', generator=g) * 0.3 for sp in specs]
    ms = [torch.randn(sp.n_params, device='cuda', generator=g) * 0.1 for sp in specs]
    vs = [torch.rand(sp.n_params, device='cuda', generator=g) * 0.01 for sp in specs]
    grads = [[torch.randn(sp.n_params, device='cuda', generator=g) for _ in range(world)] for sp in specs]
    ref = []
    for j, sp in enumerate(specs):
        gs = grads[j][0].clone()
        for w in range(1, world):
            gs += grads[j][w]
        p, m, v = ps[j].clone(), ms[j].clone(), vs[j].clone()
        F.adam_step(p, m, v, gs, 5, 1e-2, weight_decay=1e-4, grad_scale=1.0 / world)
        ref.append((p, m, v, F.mlp_pack(p, sp)))
    blobs = [torch.zeros(sp.blob_bytes, dtype=torch.uint8, device='cuda') for sp in specs]
    srcs = [grads[j][0] if world == 1 else [t.data_ptr() for t in grads[j]] for j in range(2)]
    opt = F.MlpAdamPack(specs, ps, ms, vs, srcs, blobs)
    opt.step(5, 1e-2, weight_decay=1e-4, grad_scale=1.0 / world, zero_grad=(world == 1))
    torch.cuda.synchronize()
    for j in range(2):
        assert torch.equal(ps[j], ref[j][0]) and torch.equal(ms[j], ref[j][1]) and torch.equal(vs[j], ref[j][2])
        assert torch.equal(blobs[j], ref[j][3])
        if world == 1:
            assert float(grads[j][0].abs().max()) == 0.0


def test_adam_step_peer_rejects_bad_arguments():
    from arcnerf_b200.peer import adam_step_peer
    x = torch.zeros(64, device='cuda')
    with pytest.raises(RuntimeError, match='range'):
        adam_step_peer(_ptrs([x]), None, 1, 2, 64, x, x, x, 1, 1e-3)
    with pytest.raises(RuntimeError, match='world'):
        adam_step_peer(_ptrs([x] * 17), None, 17, 0, 64, x, x, x, 1, 1e-3)


def test_peer_barrier_single_rank_and_timeout():
    """world 1 passes at once; a missing peer trips the timeout flag instead of spinning forever"""
    from arcnerf_b200 import _lib
    flags = torch.zeros(32, dtype=torch.int64, device='cuda')
    epoch = torch.zeros(1, dtype=torch.int64, device='cuda')
    status = torch.zeros(1, dtype=torch.int32, device='cuda')
    fp = _ptrs([flags])
    for _ in range(3):
        _lib.call('arcb200_peer_barrier', fp, 1, 0, epoch.data_ptr(), status.data_ptr(), 100, _lib.current_stream())
    torch.cuda.synchronize()
    assert int(epoch) == 3 and int(flags[0]) == 3 and int(status) == 0
    # two 'ranks' sharing one flag array layout, but rank 1 never arrives
    other = torch.zeros(32, dtype=torch.int64, device='cuda')
    fp2 = _ptrs([flags, other])
    _lib.call('arcb200_peer_barrier', fp2, 2, 0, epoch.data_ptr(), status.data_ptr(), -50, _lib.current_stream())   # < 0: probing mode, no trap
    torch.cuda.synchronize()
    assert int(status) == 1 and int(other[0]) == 4


def _two_rank_worker(rank, world, port, out_dir, one_device):
    """one_device: both ranks on GPU 0 (a 1-GPU lease).  Two processes can IPC-map each other's arenas on one device, so the
    arena mapping and both peer exchange kernels (loads from / stores to the other process's arena) run as on two GPUs.  Two
    things differ there: NCCL refuses two ranks on one device, so the process group (handle exchange, the all-reduce of the
    checked alternative) is gloo; and the GPU time-slices the two processes, so a kernel spinning on the other rank's flag is
    not guaranteed to make progress - the flag barrier is replaced by stream-synchronise + host barrier
    (ARCB200_PEER_HOST_BARRIER).  The flag-barrier kernel itself is covered by test_peer_barrier_single_rank_and_timeout and,
    on a box with two GPUs, by this test."""
    import torch.distributed as dist
    os.environ['MASTER_ADDR'], os.environ['MASTER_PORT'] = '127.0.0.1', str(port)
    os.environ.setdefault('ARCB200_PEER_TIMEOUT_MS', '30000')       # a stuck barrier fails the test instead of hanging it
    dev = 0 if one_device else rank
    if one_device:
        os.environ['ARCB200_PEER_HOST_BARRIER'] = '1'
    torch.cuda.set_device(dev)
    if one_device:
        dist.init_process_group('gloo', rank=rank, world_size=world)
    else:
        dist.init_process_group('nccl', rank=rank, world_size=world, device_id=torch.device('cuda', rank))
    from arcnerf_b200 import synth
    from arcnerf_b200.ngp import NGP_DEFAULTS, NGPPipeline, init_params
    params = init_params(dict(NGP_DEFAULTS), seed=3, grid_scale=0.5)
    n = 1024
    res = {}
    # 'mc16' (in-switch reduce) needs two real GPUs behind an NVSwitch; two ranks on one device cannot form a multicast group
    peer_modes = ('peer', 'peer16') if one_device else ('peer', 'peer16', 'mc16')
    for mode in ('nccl', ) + peer_modes:
        pipe = NGPPipeline(max_rays=n, capacity=1 << 17, device='cuda:%d' % dev, world_size=world, exchange=mode,
                           params={k: v.clone() for k, v in params.items()}, lr=1e-2)
        pipe.set_occupancy(synth.occupancy_volume('lego'))
        losses = []
        for step in range(4):
            o, d = synth.train_ray_batch(n, seed=100 + step * world + rank)
            t = np.random.RandomState(500 + step * world + rank).uniform(0, 1, (n, 3)).astype(np.float32)
            dv = [torch.from_numpy(x).cuda() for x in (o, d, t)]
            losses.append(float(pipe.train_step(*dv)))
        torch.cuda.synchronize()
        if mode != 'nccl':
            assert pipe.arena is not None and pipe.exchange == mode, 'peer arena was not mapped: fell back to %s' % pipe.exchange
            pipe.arena.check()
            pipe_table = pipe.gather_master_table().clone()
        else:
            pipe_table = pipe.table.clone()
        res[mode] = dict(table=pipe_table, table_h=pipe.table_h.clone(), geo=pipe.geo_p.clone(), rad=pipe.rad_p.clone(), loss=losses)
        pipe.close()                                   # collective teardown of the arena before the next mode allocates one
    ok = True
    msg = []
    for mode_b in peer_modes:
        for k in ('table', 'table_h', 'geo', 'rad'):
            # the hash-grid backward accumulates with float atomics (run-to-run order differs), and Adam turns a sign flip
            # of a near-zero gradient into a +-lr step: compare by the fraction of entries that moved apart
            a, b = res['nccl'][k].float(), res[mode_b][k].float()
            far = float(((a - b).abs() > 2e-3).float().mean())
            if far > 2e-3:
                ok = False
                msg.append('%s %s differs: fraction beyond 2e-3 = %.3e, max |d| %.3e' % (mode_b, k, far, float((a - b).abs().max())))
        if not np.allclose(res['nccl']['loss'], res[mode_b]['loss'], rtol=2e-3):
            ok = False
            msg.append('%s losses differ %s %s' % (mode_b, res['nccl']['loss'], res[mode_b]['loss']))
        # replicas agree across ranks
        th = res[mode_b]['table_h'].float()
        th = th.cpu() if one_device else th                 # gloo gathers host tensors, NCCL device tensors
        other = [torch.empty_like(th) for _ in range(world)]
        dist.all_gather(other, th)
        if not all(torch.equal(o, th) for o in other):
            ok = False
            msg.append('%s fp16 tables differ between ranks' % mode_b)
    with open(os.path.join(out_dir, 'rank%d.txt' % rank), 'w') as f:
        f.write('OK' if ok else 'FAIL ' + '; '.join(msg))
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_peer_exchange_matches_all_reduce(tmp_path):
    """two processes, IPC-mapped arenas, flag barriers, 'peer' and the default 'peer16' exchange against the all-reduce path;
    on a 1-GPU lease both ranks share device 0"""
    import torch.multiprocessing as mp
    port = 29500 + (os.getpid() % 400)
    one_device = torch.cuda.device_count() < 2
    mp.spawn(_two_rank_worker, args=(2, port, str(tmp_path), one_device), nprocs=2, join=True)
    for r in range(2):
        assert open(tmp_path / ('rank%d.txt' % r)).read() == 'OK'


@pytest.mark.parametrize('world', [2, 3, 8])
def test_half_precision_exchange_kernels(world):
    """grad_to_half + adam_step_peer16 on emulated ranks (buffers on one device) against the fp32 sum-then-Adam path fed
    with the same fp16-rounded peer gradients"""
    from arcnerf_b200 import functional as F
    from arcnerf_b200.peer import adam_step_peer16, grad_to_half, slice_bounds
    n = 40000 + 8 * 3
    g = torch.Generator(device='cuda').manual_seed(world)
    grads = [torch.randn(n, device='cuda', generator=g) * 1e-3 for _ in range(world)]
    grads[0][5] = 1e9                                             # saturates instead of overflowing
    p0 = torch.randn(n, device='cuda', generator=g)
    m0 = torch.randn(n, device='cuda', generator=g) * 0.1
    v0 = torch.rand(n, device='cuda', generator=g) * 0.01
    scale = 128.0
    bounds = [slice_bounds(n, r, world, quantum=1024) for r in range(world)]
    g16 = [torch.full((n, ), 3.0, dtype=torch.float16, device='cuda') for _ in range(world)]
    g32 = [x.clone() for x in grads]
    for r in range(world):
        grad_to_half(g32[r], g16[r], bounds[r][0], bounds[r][1], scale)
    torch.cuda.synchronize()
    for r in range(world):
        lo, hi = bounds[r]
        assert torch.equal(g32[r][lo:hi], grads[r][lo:hi]) and float(g32[r][:lo].abs().sum() + g32[r][hi:].abs().sum()) == 0.0
        ref16 = (grads[r] * scale).clamp(-65504, 65504).half()
        assert torch.equal(g16[r][:lo], ref16[:lo]) and torch.equal(g16[r][hi:], ref16[hi:])
        assert bool((g16[r][lo:hi] == 3.0).all())                 # the own slice is not converted
    halves = [torch.full((n, ), 7.0, dtype=torch.float16, device='cuda') for _ in range(world)]
    p_all, h_ref = p0.clone(), torch.empty(n, dtype=torch.float16, device='cuda')
    for r in range(world):
        lo, hi = bounds[r]
        if hi == lo:
            continue
        p, m, v = p0.clone(), m0.clone(), v0.clone()
        adam_step_peer16(g32[r], _ptrs(g16), _ptrs(halves), world, r, lo, hi, p, m, v, 2, 1e-2, scale=scale, grad_scale=1.0 / world)
        # reference: own fp32 + peers' fp16 values (summed in rank order, then / scale) -> single-GPU Adam
        s = torch.zeros(hi - lo, device='cuda')
        for w in range(world):
            if w != r:
                s += g16[w][lo:hi].float()
        gsum = grads[r][lo:hi] + s * (1.0 / scale)
        pr, mr, vr = p0[lo:hi].clone(), m0[lo:hi].clone(), v0[lo:hi].clone()
        hr = torch.empty(hi - lo, dtype=torch.float16, device='cuda')
        F.adam_step(pr, mr, vr, gsum, 2, 1e-2, p_half=hr, grad_scale=1.0 / world, zero_grad=False)
        torch.cuda.synchronize()
        assert torch.equal(p[lo:hi], pr) and torch.equal(m[lo:hi], mr) and torch.equal(v[lo:hi], vr)
        assert torch.equal(p[:lo], p0[:lo]) and torch.equal(p[hi:], p0[hi:])
        assert float(g32[r][lo:hi].abs().sum()) == 0.0            # own slice cleared after use
        h_ref[lo:hi] = hr
    for h in halves:
        assert torch.equal(h, h_ref)
