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
