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
W, E, D, K, BUCKET = 40, 1024, 4, 11, 4  # K is not a multiple of the bucket: the partial block is flushed on read


def _worker(rank: int, world_size: int, port: int, out_dir: str, dag: str) -> None:
    import torch
    import torch.distributed as dist

    from fimo_b200 import library

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
            rig.step(K)
            out = rig.download()
            peers = np.stack([sharding.global_stats(lib, rig.group.handle, g, abi.STATS_MODE_PEERS) for g in range(K - 8, K)])
            nccl = np.stack([sharding.global_stats(lib, rig.group.handle, g, abi.STATS_MODE_NCCL) for g in range(K - 3, K)])
            dist.barrier()  # nobody tears its inbox down while a peer may still store into it
        finally:
            rig.close()
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), peers=peers, nccl=nccl, **{k: v for k, v in out.items() if k not in synth.BATCH_SCOPE})
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("dag", ["chain3", "dag16"])
def test_two_gpus_equal_one_batch_on_the_oracle(oracle, tmp_path, dag):
    import torch
    import torch.multiprocessing as mp

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 CUDA devices (gpurun --gpus 2)")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(2, port, str(tmp_path), dag), nprocs=2, join=True)

    data = synth.generate(W, E, D, dag=dag)
    ref = synth.WorldStepRig(oracle, W, E, D, data=data, dag=dag)
    per_step = []
    try:
        for _ in range(K):
            ref.step(1)
            per_step.append(ref.stats().copy())
        want = ref.download()
    finally:
        ref.close()
    shards = [np.load(tmp_path / f"rank{r}.npz") for r in range(2)]
    for name in synth.outputs_for(dag):
        if name in synth.BATCH_SCOPE:
            continue
        got = np.concatenate([s[name] for s in shards], axis=0)
        assert np.array_equal(got.view(np.uint32), want[name].view(np.uint32)), name
    for s in shards:  # every rank holds the same global sums, for every generation, on both transports
        for i, g in enumerate(range(K - 8, K)):
            assert np.array_equal(s["peers"][i][:5], per_step[g][:5]), ("peers", g, s["peers"][i], per_step[g])
        for i, g in enumerate(range(K - 3, K)):
            assert np.array_equal(s["nccl"][i][:5], per_step[g][:5]), ("nccl", g, s["nccl"][i], per_step[g])
