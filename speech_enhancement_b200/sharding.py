"""Utterance sharding across the GPUs of one box.

Every op on the hot path is per-utterance, so inference shards the batch dimension contiguously
-- rank r of G takes x[r*B/G : (r+1)*B/G] (remainder to the lowest ranks) -- with NO collective on
the data path.  The only communication is optional and at the edges: gathering the enhanced
waveforms on rank 0, and reducing a timing / metric scalar as MAX / SUM over ranks.  The process
group is whatever the launcher initialised (`nccl` on GPUs; `gloo` in the CPU tests).
"""
import torch
import torch.distributed as dist


def shard_bounds(n, rank, world):
    """[begin, end) of rank `rank`'s contiguous share of n utterances."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError(f'shard_bounds: bad rank {rank} / world {world}')
    base, rem = divmod(n, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def shard_batch(x, rank=None, world=None):
    """this rank's utterances of a [B, ...] batch (a view, no copy)."""
    if rank is None:
        rank = dist.get_rank() if dist.is_initialized() else 0
    if world is None:
        world = dist.get_world_size() if dist.is_initialized() else 1
    b, e = shard_bounds(x.shape[0], rank, world)
    return x[b:e]


def gather_utterances(local, total, dst=0):
    """collect per-rank [b_r, L] results into [total, L] on rank `dst` (None elsewhere); ragged shares
    are padded to the largest share for the collective and cut back afterwards."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return local
    world, rank = dist.get_world_size(), dist.get_rank()
    sizes = [shard_bounds(total, r, world) for r in range(world)]
    most = max(e - b for b, e in sizes)
    pad = torch.zeros((most,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[:local.shape[0]] = local
    bufs = [torch.empty_like(pad) for _ in range(world)] if rank == dst else None
    dist.gather(pad, bufs, dst=dst)
    if rank != dst:
        return None
    return torch.cat([bufs[r][:e - b] for r, (b, e) in enumerate(sizes)], 0)


def reduce_max(value, device=None):
    """max over ranks of a python float (step time: the slowest rank defines the job)."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    if t.is_cuda:
        t = t.float()
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def enhance_sharded(model, x, dst=0):
    """run `model` (returns (spec, wav)) on this rank's share of x [B, L]; rank `dst` gets all waveforms."""
    mine = shard_batch(x)
    dev = next(model.parameters()).device
    with torch.no_grad():
        wav = model(mine.to(dev))[1] if mine.shape[0] else None
    if wav is None:
        # an empty share still has to take part in the gather with the right trailing shape
        probe = model(x[:1].to(dev))[1]
        wav = probe[:0]
    return gather_utterances(wav, x.shape[0], dst=dst)
