"""Host-side data-parallel plumbing (one process per GPU, torch.distributed): contiguous ray sharding for frame
rendering / voxel slabs, the per-step gradient exchange, and the final gather.  Backend-agnostic (NCCL on the GPUs,
gloo in the CPU tests).  The reference gets the gradient exchange from DDP (common/trainer/basic_trainer.py:192-202)
and shards per image with DistributedSampler (arcnerf/trainer/arcnerf_trainer.py:242-249)."""
import torch


def shard_range(n, rank, world):
    """contiguous [start, end) of rank's shard of n units; shards differ in size by at most one tile of ceil(n/world)"""
    per = (n + world - 1) // world
    return min(n, rank * per), min(n, (rank + 1) * per)


def deal_tiles(n, rank, world, tile=256):
    """Round-robin shard for frame rendering: tiles of `tile` consecutive units are dealt to the ranks in turn, so that every
    rank gets its share of the rays that hit the object (a contiguous split leaves the border ranks idle and the frame time
    is the slowest rank's).  -> (index (m,) int64 of this rank's units, per = padded share size, the same on every rank)"""
    n_tiles = (n + tile - 1) // tile
    tiles_per = (n_tiles + world - 1) // world
    idx = ((torch.arange(tiles_per)[:, None] * world + rank) * tile + torch.arange(tile)[None, :]).reshape(-1)
    return idx[idx < n], tiles_per * tile


def gather_dealt(local, n_total, world, tile=256, out=None):
    """all-gather the shares made by deal_tiles of an (n_total, C) tensor and put the tiles back in order on EVERY rank:
    one collective into a preallocated buffer + one strided copy.  `local`: this rank's (m, C) rows in deal order."""
    import torch.distributed as dist
    if world == 1:
        return local
    n_tiles = (n_total + tile - 1) // tile
    tiles_per = (n_tiles + world - 1) // world
    c = local.shape[1]
    share = torch.zeros((tiles_per * tile, c), dtype=local.dtype, device=local.device)
    share[:local.shape[0]] = local
    gathered = torch.empty((world, tiles_per, tile, c), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(gathered.view(-1, c), share)
    if out is None:
        out = torch.empty((tiles_per, world, tile, c), dtype=local.dtype, device=local.device)
    out.view(tiles_per, world, tile, c).copy_(gathered.permute(1, 0, 2, 3))
    return out.view(-1, c)[:n_total]


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
