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


def bind_near_gpu(local_rank: int) -> list:
    """Pin this process to the CPUs the driver reports as closest to GPU ``local_rank``.

    With one process per GPU on a multi-socket box, page-locked result buffers then come from the
    memory of the GPU's own socket and device-to-host copies do not cross the socket link.
    Returns the CPU list (empty when NVML or the affinity call is unavailable; nothing changes
    then).  ``RSB_NO_BIND=1`` turns it off."""
    if os.environ.get("RSB_NO_BIND") == "1" or not hasattr(os, "sched_setaffinity"):
        return []
    try:
        import pynvml

        pynvml.nvmlInit()
        handle = pynvml.nvmlDeviceGetHandleByIndex(int(local_rank))
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(handle, words)
        cpus = [64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1]
        allowed = os.sched_getaffinity(0)
        cpus = [c for c in cpus if c in allowed]
        if cpus:
            os.sched_setaffinity(0, cpus)
        return cpus
    except Exception:
        return []
