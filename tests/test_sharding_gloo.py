"""The N>1 host logic on CPU: world_size-2 gloo, worlds sharded by global index, stats all-reduced.
Each rank steps its shard on the CPU oracle; the gathered result must equal a single-process run bit for bit
(the generator is counter-based and worlds are independent)."""
import os
import socket

import numpy as np
import pytest

from fimo_b200 import abi, sharding, synth

W, E, D, K = 6, 128, 4, 3


def _worker(rank: int, world_size: int, port: int, out_dir: str) -> None:
    import torch
    import torch.distributed as dist

    from oracle import loader

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    try:
        first, count = sharding.shard_range(rank, world_size, W)
        rig = synth.WorldStepRig(loader.load(), count, E, D, world0=first)
        try:
            rig.step(K)
            out = rig.download()
        finally:
            rig.close()
        total = sharding.allreduce_stats(dist, torch, out["stats"].reshape(-1).view(np.uint64))
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), first=first, count=count, total=total,
                 **{k: v for k, v in out.items() if k != "stats"})
    finally:
        dist.destroy_process_group()


def test_shard_range_partitions_all_worlds():
    for n, total in ((1, 7), (2, 7), (4, 4096), (8, 32768), (8, 5)):
        spans = [sharding.shard_range(r, n, total) for r in range(n)]
        assert spans[0][0] == 0 and sum(c for _, c in spans) == total
        for (f0, c0), (f1, _) in zip(spans, spans[1:]):
            assert f0 + c0 == f1


def test_two_ranks_reproduce_the_single_process_result(oracle, tmp_path):
    import torch.multiprocessing as mp

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)

    ref = synth.WorldStepRig(oracle, W, E, D)
    try:
        ref.step(K)
        want = ref.download()
    finally:
        ref.close()
    shards = [np.load(tmp_path / f"rank{r}.npz") for r in range(2)]
    for name in ("pos", "world", "vis_mask", "vis_list", "vis_count"):
        got = np.concatenate([s[name] for s in shards], axis=0)
        assert np.array_equal(got.view(np.uint32), want[name].view(np.uint32)), name
    st = want["stats"].reshape(-1).view(np.uint64)
    for s in shards:  # every rank holds the same global sums
        for word in (abi.STAT_VISIBLE, abi.STAT_ENTITIES, abi.STAT_MASK_SUM, abi.STAT_WORLDS, abi.STAT_STEPS):
            assert s["total"][word] == st[word]
