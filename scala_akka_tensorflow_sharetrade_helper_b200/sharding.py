"""Episode sharding for the multi-GPU path: contiguous blocks of episodes per rank,
so a ticker's episodes stay on one GPU; the RNG counter of an episode is its GLOBAL index, so a
sharded run draws exactly the numbers of the single-GPU run."""
from __future__ import annotations


def shard(n_total: int, rank: int, world: int) -> tuple[int, int]:
    """(first global episode, number of episodes) of `rank`; remainders go to the low ranks."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("rank/world out of range")
    base, rem = divmod(int(n_total), int(world))
    first = rank * base + min(rank, rem)
    return first, base + (1 if rank < rem else 0)
