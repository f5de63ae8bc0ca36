"""World-index sharding across ranks: one process per GPU, worlds are independent, so the
only exchange is the end-of-step statistics all-reduce.  Used by bench.py (NCCL) and tests (gloo on CPU)."""
from __future__ import annotations

import numpy as np

from . import abi


def shard_range(rank: int, world_size: int, total_worlds: int) -> tuple[int, int]:
    """Contiguous block [first, first + count) of global world indices owned by `rank`; the first
    `total % world_size` ranks own one extra world."""
    base, extra = divmod(total_worlds, world_size)
    first = rank * base + min(rank, extra)
    return first, base + (1 if rank < extra else 0)


def allreduce_stats(dist, torch, stats_u64: np.ndarray, device=None) -> np.ndarray:
    """Sum of the per-rank FIMO_B200_STAT_* words over all ranks (every word is an additive count)."""
    t = torch.from_numpy(stats_u64[:abi.STAT_WORDS].astype(np.int64))
    if device is not None:
        t = t.to(device)
    dist.all_reduce(t)
    return t.cpu().numpy().astype(np.uint64)
