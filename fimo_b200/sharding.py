"""World-index sharding across ranks: one process per GPU, worlds are independent units
(modules/fimo_worlds/src/World.zig:26,45-52), so the only exchange is the end-of-step statistics.

The exchange itself is product code (csrc/stats_exchange.cu + the publishing CTA of the statistics kernel); this
module is the harness side of it: which worlds a rank owns, and the out-of-band hand-over of the 64-byte peer
handles / the 128-byte NCCL id through `torch.distributed` (any backend: NCCL on GPUs, gloo in the CPU tests).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import abi


def shard_range(rank: int, world_size: int, total_worlds: int) -> tuple[int, int]:
    """Contiguous block [first, first + count) of global world indices owned by `rank`; the first
    `total % world_size` ranks own one extra world."""
    base, extra = divmod(total_worlds, world_size)
    first = rank * base + min(rank, extra)
    return first, base + (1 if rank < extra else 0)


def allreduce_stats(dist, torch, stats_u64: np.ndarray, device=None) -> np.ndarray:
    """Sum of the per-rank FIMO_B200_STAT_* words over all ranks with the harness's own collective (every word is
    an additive count) — the independent check of the product's exchange."""
    t = torch.from_numpy(stats_u64[:abi.STAT_WORDS].astype(np.int64))
    if device is not None:
        t = t.to(device)
    dist.all_reduce(t)
    return t.cpu().numpy().astype(np.uint64)


def _all_gather_bytes(dist, torch, payload: bytes, device=None) -> list[bytes]:
    mine = torch.frombuffer(bytearray(payload), dtype=torch.uint8)
    if device is not None:
        mine = mine.to(device)
    parts = [torch.empty_like(mine) for _ in range(dist.get_world_size())]
    dist.all_gather(parts, mine)
    return [bytes(p.cpu().numpy().tobytes()) for p in parts]


def bind_stats_peers(lib: abi.WorldsLib, group_handle, dist, torch, slots: int = 64, device=None) -> None:
    """fimo_b200_group_stats_peers_open on every rank, all-gather of the handles, fimo_b200_group_stats_peers_bind."""
    handle = (C.c_uint8 * abi.STATS_PEER_HANDLE_BYTES)()
    lib.check(lib.fimo_b200_group_stats_peers_open(group_handle, slots, handle))
    handles = b"".join(_all_gather_bytes(dist, torch, bytes(handle), device))
    buf = (C.c_uint8 * len(handles)).from_buffer_copy(handles)
    lib.check(lib.fimo_b200_group_stats_peers_bind(group_handle, buf, dist.get_rank(), dist.get_world_size()))


def bind_stats_nccl(lib: abi.WorldsLib, group_handle, dist, torch, bucket: int = 8, device=None) -> None:
    """ncclGetUniqueId on rank 0, hand-over to every rank, fimo_b200_group_stats_nccl_bind (collective)."""
    uid = (C.c_uint8 * abi.STATS_NCCL_ID_BYTES)()
    if dist.get_rank() == 0:
        lib.check(lib.fimo_b200_stats_nccl_unique_id(uid))
    uid0 = _all_gather_bytes(dist, torch, bytes(uid), device)[0]
    buf = (C.c_uint8 * len(uid0)).from_buffer_copy(uid0)
    lib.check(lib.fimo_b200_group_stats_nccl_bind(group_handle, buf, dist.get_rank(), dist.get_world_size(), bucket))


def global_stats(lib: abi.WorldsLib, group_handle, generation: int, mode: int) -> np.ndarray:
    """The statistics of `generation` summed over all ranks by the product's exchange (8 u64 words)."""
    out = (C.c_uint64 * abi.STAT_WORDS)()
    lib.check(lib.fimo_b200_group_global_stats(group_handle, generation, mode, out))
    return np.array(list(out), dtype=np.uint64)
