"""Utterance sharding for multi-GPU synthesis.

The path has no cross-utterance dependency, so the batch is split contiguously over ranks
(one process per GPU) and no collective is needed on the data path; weights are replicated.
"""
from __future__ import annotations


def shard_bounds(n_items: int, world_size: int, rank: int) -> tuple[int, int]:
    """[lo, hi) of rank's contiguous shard; the first n_items % world_size ranks get one extra."""
    if world_size <= 0 or not (0 <= rank < world_size):
        raise ValueError(f"bad rank {rank} / world_size {world_size}")
    base, rem = divmod(n_items, world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_batch(tensor, world_size: int, rank: int):
    """Rank's slice along dim 0 (a view)."""
    lo, hi = shard_bounds(tensor.shape[0], world_size, rank)
    return tensor[lo:hi]


def max_over_ranks(values, device=None):
    """Element-wise MAX over all ranks of a list of floats (per-rank timings -> the job's time).  The only collective
    of this path (bench.py; NCCL on the GPU box, gloo in the CPU tests); identity without a process group."""
    import torch
    import torch.distributed as dist
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(x) for x in t]


def gather_shards(local, n_items: int, world_size: int, rank: int):
    """Optional assembly of per-rank result shards [n_local, ...] into [n_items, ...] on every rank (SURVEY.md section 8e:
    "optional all_gather of the wav shards"); ragged shards are padded to the largest for the collective."""
    import torch
    import torch.distributed as dist
    if world_size == 1:
        return local
    sizes = [shard_bounds(n_items, world_size, r) for r in range(world_size)]
    biggest = max(hi - lo for lo, hi in sizes)
    pad = torch.zeros((biggest,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    parts = [torch.empty_like(pad) for _ in range(world_size)]
    dist.all_gather(parts, pad)
    return torch.cat([p[: hi - lo] for p, (lo, hi) in zip(parts, sizes)])
