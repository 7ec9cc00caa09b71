"""Batch sharder: independent images are split across the GPUs of one box, one process per
GPU, no collective on the data path.  Rank r uses seed ``seed + r`` (so its shard equals the
single-process run with that seed) and writes its own ``.h5`` shards."""

from __future__ import annotations

import os

__all__ = ["rank_info", "rank_seed", "shard_bundles", "shard_name"]


def rank_info():
    """(rank, world, local_rank) from the torchrun environment (1 process when absent)."""
    return (int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)),
            int(os.environ.get("LOCAL_RANK", 0)))


def rank_seed(seed, rank):
    return None if seed is None else int(seed) + int(rank)


def shard_bundles(num_bundles: int, rank: int, world: int) -> range:
    """Contiguous, balanced range of bundle indices owned by ``rank``."""
    base, extra = divmod(num_bundles, world)
    start = rank * base + min(rank, extra)
    return range(start, start + base + (1 if rank < extra else 0))


def shard_name(mode: str, opt: str, index: int, rank: int, world: int) -> str:
    """The reference overwrites ``data_{mode}_{opt}.h5`` for every bundle
    (``simulations.py:91``); shards here carry the bundle index (and the rank when sharded)."""
    if world > 1:
        return f"data_{mode}_{opt}_r{rank}_{index}.h5"
    return f"data_{mode}_{opt}_{index}.h5"
