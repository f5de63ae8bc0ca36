"""N > 1 on real GPUs (-m gpu; skipped on a box with fewer than 2 devices): worlds sharded by global index over two
processes / two GPUs, statistics exchanged by the product (peer stores over NVLink AND the NCCL all-reduce, both bound
on the same group).  The concatenated shards equal the oracle's single-batch run bit for bit, and the exchanged
statistics of every generation equal the oracle's (worlds are independent units, World.zig:26,45-52)."""
import os
import socket

import numpy as np
import pytest

from fimo_b200 import abi, sharding, synth

pytestmark = pytest.mark.gpu
D, BUCKET = 4, 4
# (worlds, entities, queued generations).  K is never a multiple of the bucket: the partial block is flushed on read.
# SMALL stays inside the NCCL ring (4 x BUCKET = 16 rows) and every generation is its own launch.  LONG has enough worlds
# per GPU for queued generations to share launches (up to 16 each, test_gpu_parity.py QUEUED_CASES) and wraps the ring
# three times: a launch must never write ring rows whose generations still wait for their all-reduce.
SMALL, LONG = (40, 1024, 11), (2400, 2048, 53)
NCCL_READABLE = 3 * BUCKET  # rows of the ring that are never about to be overwritten


def _worker(rank: int, world_size: int, port: int, out_dir: str, dag: str, shape) -> None:
    import torch
    import torch.distributed as dist

    from fimo_b200 import library

    W, E, K = shape
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)  # handles / ids travel out of band; any backend does
    lib = library.load()
    try:
        first, count = sharding.shard_range(rank, world_size, W)
        rig = synth.WorldStepRig(lib, count, E, D, world0=first, options=synth.RigOptions(device=rank), dag=dag)
        try:
            sharding.bind_stats_peers(lib, rig.group.handle, dist, torch, slots=8)
            sharding.bind_stats_nccl(lib, rig.group.handle, dist, torch, bucket=BUCKET)
            before = rig.kernels_launched()
            rig.step(2)  # two generations first: the shared launches that follow start in the middle of a bucket
            rig.step(K - 2)
            launches = rig.kernels_launched() - before
            out = rig.download()
            peers = np.stack([sharding.global_stats(lib, rig.group.handle, g, abi.STATS_MODE_PEERS) for g in range(K - 8, K)])
            nccl = np.stack([sharding.global_stats(lib, rig.group.handle, g, abi.STATS_MODE_NCCL) for g in range(max(0, K - NCCL_READABLE), K)])
            dist.barrier()  # nobody tears its inbox down while a peer may still store into it
        finally:
            rig.close()
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), peers=peers, nccl=nccl, launches=launches, **{k: v for k, v in out.items() if k not in synth.BATCH_SCOPE})
    finally:
        dist.destroy_process_group()


def _oracle_statistics(oracle, dag, shape):
    W, E, K = shape
    data = synth.generate(W, E, D, dag=dag)
    ref = synth.WorldStepRig(oracle, W, E, D, data=data, dag=dag)
    per_step = []
    try:
        for _ in range(K):
            ref.step(1)
            per_step.append(ref.stats().copy())
        return per_step, ref.download()
    finally:
        ref.close()


@pytest.mark.parametrize("shape", [SMALL, (LONG[0] // 2, LONG[1], LONG[2])], ids=["small", "long"])
def test_nccl_ring_of_one_rank_holds_every_readable_generation(oracle, product, monkeypatch, shape):
    """The NCCL transport on ONE GPU (a communicator of one rank: the all-reduce is the identity, everything around it —
    ring rows, bucket boundaries, the events between the group's stream and the collective stream, the flush of the
    partial block on read — is the code every rank runs).  K generations are queued without waiting, so they leave in
    launches of several generations ("long"); a launch must never write ring rows whose generations have not been
    reduced yet.  Every generation still inside the ring equals the oracle's statistics of that step.  (What one rank
    cannot show is a row reduced twice or not at all — the 2-GPU test below does.)"""
    import ctypes as C

    import torch  # noqa: F401  (brings libnccl.so.2 into the process)

    W, E, K = shape
    monkeypatch.setenv("FIMO_B200_TEAM", "1")
    per_step, want = _oracle_statistics(oracle, "chain3", shape)
    rig = synth.WorldStepRig(product, W, E, D, dag="chain3")
    try:
        uid = (C.c_uint8 * abi.STATS_NCCL_ID_BYTES)()
        product.check(product.fimo_b200_stats_nccl_unique_id(uid))
        product.check(product.fimo_b200_group_stats_nccl_bind(rig.group.handle, uid, 0, 1, BUCKET))
        before = rig.kernels_launched()
        rig.schedule(K)
        rig.wait()
        if shape != SMALL:
            assert rig.kernels_launched() - before < K  # generations did share launches
        for g in range(max(0, K - NCCL_READABLE), K):
            got = sharding.global_stats(product, rig.group.handle, g, abi.STATS_MODE_NCCL)
            assert np.array_equal(got[:5], per_step[g][:5]), (g, got, per_step[g])
        got = rig.download()
    finally:
        rig.close()
    problems = synth.compare_outputs(got, want, 0.0)
    assert not problems, "\n".join(problems)


@pytest.mark.parametrize("dag,shape", [("chain3", SMALL), ("dag16", SMALL), ("chain3", LONG)], ids=["chain3-small", "dag16-small", "chain3-long"])
def test_two_gpus_equal_one_batch_on_the_oracle(oracle, tmp_path, monkeypatch, dag, shape):
    import torch
    import torch.multiprocessing as mp

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 CUDA devices (gpurun --gpus 2)")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    W, E, K = shape
    monkeypatch.setenv("FIMO_B200_TEAM", "1")  # inherited by the workers: one CTA per world, the plan that shares launches
    if shape == LONG:  # ... and every launch takes as many generations as the plan allows, whatever the host's pace
        monkeypatch.setenv("FIMO_B200_GENS_HOLD", "16")
    mp.spawn(_worker, args=(2, port, str(tmp_path), dag, shape), nprocs=2, join=True)

    per_step, want = _oracle_statistics(oracle, dag, shape)
    shards = [np.load(tmp_path / f"rank{r}.npz") for r in range(2)]
    if shape == LONG:
        assert all(int(s["launches"]) < K for s in shards)  # generations did share launches
    for name in synth.outputs_for(dag):
        if name in synth.BATCH_SCOPE:
            continue
        got = np.concatenate([s[name] for s in shards], axis=0)
        assert np.array_equal(got.view(np.uint32), want[name].view(np.uint32)), name
    for s in shards:  # every rank holds the same global sums, for every generation, on both transports
        for i, g in enumerate(range(K - 8, K)):
            assert np.array_equal(s["peers"][i][:5], per_step[g][:5]), ("peers", g, s["peers"][i], per_step[g])
        for i, g in enumerate(range(max(0, K - NCCL_READABLE), K)):
            assert np.array_equal(s["nccl"][i][:5], per_step[g][:5]), ("nccl", g, s["nccl"][i], per_step[g])
