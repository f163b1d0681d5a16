"""Host-side data-parallel plumbing (one process per GPU, torch.distributed): contiguous ray sharding for frame
rendering / voxel slabs, the per-step gradient exchange, and the final gather.  Backend-agnostic (NCCL on the GPUs,
gloo in the CPU tests).  The reference gets the gradient exchange from DDP (common/trainer/basic_trainer.py:192-202)
and shards per image with DistributedSampler (arcnerf/trainer/arcnerf_trainer.py:242-249)."""
import torch


def shard_range(n, rank, world):
    """contiguous [start, end) of rank's shard of n units; shards differ in size by at most one tile of ceil(n/world)"""
    per = (n + world - 1) // world
    return min(n, rank * per), min(n, (rank + 1) * per)


def all_reduce_sum_(tensors):
    """in-place sum over ranks of every tensor in the list (hash-table gradient, MLP gradients)"""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return tensors
    for t in tensors:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return tensors


def gather_rows(local, n_total, dst=0):
    """gather row shards (made by shard_range) of an (n_total, ...) tensor to rank dst; returns the full tensor there"""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return local
    world, rank = dist.get_world_size(), dist.get_rank()
    per = (n_total + world - 1) // world
    pad = torch.zeros((per,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[:local.shape[0]] = local
    outs = [torch.empty_like(pad) for _ in range(world)] if rank == dst else None
    dist.gather(pad, outs, dst=dst)
    if rank != dst:
        return None
    return torch.cat(outs, 0)[:n_total]
