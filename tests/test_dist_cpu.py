"""CPU (gloo, world size 2): the multi-GPU host logic - ray sharding, gradient exchange + 1/world scaling, final gather."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from arcnerf_b200.dist import all_reduce_sum_, deal_tiles, gather_dealt, gather_rows, shard_range


def _free_port():
    s = socket.socket(); s.bind(('127.0.0.1', 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, n_total, q):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        # per-rank "gradients": what the backward kernels would leave in the flat buffers
        g_table = torch.arange(1000, dtype=torch.float32) * (rank + 1)
        g_mlp = torch.ones(77) * (10 ** rank)
        all_reduce_sum_([g_table, g_mlp])
        ok = torch.equal(g_table, torch.arange(1000, dtype=torch.float32) * 3) and torch.equal(g_mlp, torch.ones(77) * 11)
        # Adam consumes grad * (1/world): the mean over ranks, as DDP delivers
        mean = g_table * (1.0 / world)
        ok = ok and torch.allclose(mean, torch.arange(1000, dtype=torch.float32) * 1.5)
        # frame rendering: contiguous shards, gathered back in order on rank 0
        s, e = shard_range(n_total, rank, world)
        full = torch.arange(n_total * 3, dtype=torch.float32).view(n_total, 3)
        out = gather_rows(full[s:e].clone(), n_total, dst=0)
        if rank == 0:
            ok = ok and torch.equal(out, full)
        else:
            ok = ok and out is None
        # round-robin tiles (frame rendering with balanced load), reassembled on every rank
        idx, per = deal_tiles(n_total, rank, world, tile=16)
        back = gather_dealt(full[idx].clone(), n_total, world, tile=16)
        ok = ok and torch.equal(back, full) and per % 16 == 0
        q.put((rank, bool(ok), (s, e)))
    finally:
        dist.destroy_process_group()


def test_shard_range_partitions():
    for n in (0, 1, 7, 640000, 2073600):
        for world in (1, 2, 4, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert all(0 <= e - s <= (n + world - 1) // world for s, e in spans)


def test_deal_tiles_partitions():
    for n in (1, 255, 256, 1000, 640000, 2073600):
        for world in (1, 2, 3, 8):
            parts = [deal_tiles(n, r, world)[0] for r in range(world)]
            allidx = torch.cat(parts)
            assert allidx.numel() == n and torch.equal(torch.sort(allidx)[0], torch.arange(n))
            pers = {deal_tiles(n, r, world)[1] for r in range(world)}
            assert len(pers) == 1 and max(p.numel() for p in parts) <= next(iter(pers))
            # balance: no rank has more than one tile more than another
            sizes = [p.numel() for p in parts]
            assert max(sizes) - min(sizes) <= 256


def test_two_rank_exchange_and_gather():
    world, n_total = 2, 1001
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_total, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in res)
    spans = dict((r, sp) for r, _, sp in res)
    assert spans[0] == (0, 501) and spans[1] == (501, 1001)
