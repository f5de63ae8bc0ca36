"""Parity of the CUDA path with the CPU oracle, through the C ABI, on a B200 (-m gpu).

Bar (BASELINE.json north_star): integer / index / cull-mask outputs bit-exact; float transforms within
1e-5 relative.  The kernels are written so that floats are bit-exact as well (FLOAT_RTOL = 0.0 below
asserts exactly that; relax to 1e-5 only with a written reason)."""
import numpy as np
import pytest

from fimo_b200 import abi, synth

pytestmark = pytest.mark.gpu
FLOAT_RTOL = 0.0  # bit-exact; north_star tolerance is 1e-5 relative


def run_both(oracle, product, W, E, D, steps, options=None, world0=0, dag="chain3"):
    data = synth.generate(W, E, D, world0=world0, dag=dag)
    ref = synth.WorldStepRig(oracle, W, E, D, data={k: v.copy() for k, v in data.items()}, dag=dag)
    try:
        ref.step(steps)
        want = ref.download()
    finally:
        ref.close()
    dev = synth.WorldStepRig(product, W, E, D, data={k: v.copy() for k, v in data.items()}, options=options, dag=dag)
    try:
        info = dev.executor_info()
        dev.step(steps)
        got = dev.download()
    finally:
        dev.close()
    return got, want, info


CASES = [
    # (worlds, entities, depth, steps)
    (1, 1024, 4, 16),     # BASELINE configs[0]: single world, 1k entities
    (64, 1024, 4, 16),    # configs[1] shape at a size the oracle finishes in seconds
    (7, 1000, 4, 5),      # ragged: entities not a multiple of 32 / 4
    (3, 33, 4, 3),
    (5, 1, 4, 2),         # one entity per world
    (2, 4096, 4, 4),      # configs[3] world shape
    (300, 96, 8, 3),      # more worlds than resident CTAs, depth 8
]


@pytest.mark.parametrize("W,E,D,steps", CASES)
@pytest.mark.parametrize("mode", ["fused", "fused-one-cta-per-world", "unfused-graph", "unfused-streams"])
def test_three_system_chain_matches_oracle(oracle, product, monkeypatch, W, E, D, steps, mode):
    if mode == "fused-one-cta-per-world":
        monkeypatch.setenv("FIMO_B200_TEAM", "1")  # small batches default to CTA teams; pin the other kernel too
    options = synth.RigOptions(fused=mode.startswith("fused"), graph=(mode != "unfused-streams"))
    got, want, info = run_both(oracle, product, W, E, D, steps, options)
    assert info.systems == 3 and info.device_systems == 3
    if mode.startswith("fused"):
        assert info.fused_world_step == 1 and info.nodes == 1
        assert info.kernel_launches_per_generation == (1 if mode == "fused-one-cta-per-world" else info.kernel_launches_per_generation)
        assert info.kernel_launches_per_generation in (1, 2)  # 2 = team kernel + list compaction
    else:
        assert info.fused_world_step == 0 and info.nodes == 3
    assert info.uses_cuda_graph == (0 if mode == "unfused-streams" else 1)
    problems = synth.compare_outputs(got, want, FLOAT_RTOL)
    assert not problems, "\n".join(problems)
    st = got["stats"].reshape(-1).view(np.uint64)
    assert st[abi.STAT_STEPS] == W * steps and st[abi.STAT_WORLDS] == W


@pytest.mark.parametrize("fused", [True, False])
def test_empty_worlds(oracle, product, fused):
    """Worlds without entities: nothing to step, but a generation still runs — per-world counts are 0 and the batch
    statistics count the worlds and the steps, as on the oracle."""
    got, want, info = run_both(oracle, product, 3, 0, 4, 2, synth.RigOptions(fused=fused))
    assert info.systems == 3 and info.fused_world_step == 0  # nothing to fuse over
    problems = synth.compare_outputs(got, want, FLOAT_RTOL)
    assert not problems, "\n".join(problems)
    st = got["stats"].reshape(-1).view(np.uint64)
    assert st[abi.STAT_WORLDS] == 3 and st[abi.STAT_STEPS] == 6 and st[abi.STAT_VISIBLE] == 0 and st[abi.STAT_ENTITIES] == 0


TEAM_CASES = [
    # (worlds, entities, depth, steps, forced team size or None = heuristic)
    (3, 65536, 8, 3, None),    # configs[2] world shape: a world is stepped by ~197 CTAs
    (16, 1024, 4, 4, None),    # more CTAs than chunks per world
    (5, 4096, 4, 3, 3),        # team size that does not divide the chunk count
    (2, 8192, 8, 2, 37),
    (40, 2048, 8, 3, 2),
    (9, 640, 5, 4, 4),         # last chunk ragged, 5 levels
    (300, 32768, 8, 2, None),  # many LARGE worlds: the window would hold < 30 % of the parents -> teams of 16 by plan
]


@pytest.mark.parametrize("W,E,D,steps,team", TEAM_CASES)
def test_cta_teams_match_oracle(oracle, product, monkeypatch, W, E, D, steps, team):
    """Large worlds / few worlds: several CTAs step one world together (dataflow level counters through L2,
    list compaction in a second kernel).  Same bit-exact bar."""
    if team is not None:
        monkeypatch.setenv("FIMO_B200_TEAM", str(team))
    got, want, info = run_both(oracle, product, W, E, D, steps)
    assert info.fused_world_step == 1 and info.kernel_launches_per_generation == 2
    problems = synth.compare_outputs(got, want, FLOAT_RTOL)
    assert not problems, "\n".join(problems)
    st = got["stats"].reshape(-1).view(np.uint64)
    assert st[abi.STAT_STEPS] == W * steps and st[abi.STAT_WORLDS] == W and st[abi.STAT_ENTITIES] == W * E


TAIL_CASES = [
    # (dag, worlds, entities, depth, steps, FIMO_B200_PIPE or None, FIMO_B200_TAIL or None, graph)
    ("chain3", 1300, 64, 4, 3, None, "1", True),         # 2 rounds of 592 slots + 116 worlds: tail grid of 116 CTAs
    ("chain3", 1300, 64, 4, 3, None, "2", False),        # a whole round joins the tail (708 CTAs); plain stream launches
    ("chain3", 700, 520, 5, 3, "512,1,3", "1", True),    # 148 slots: 4 rounds + 108, two chunks per world (ragged)
    ("dag8", 1250, 100, 4, 3, None, None, True),          # the 8-system set uses the tail grid by default
    ("dag16", 1190, 36, 3, 4, None, "1", True),           # 2 rounds + 6 worlds: a round joins them (tail of 598)
    ("dag16", 640, 1024, 4, 2, "256,2,3", "3", False),   # 296 slots: 1 fixed round, 344 tail CTAs
]


@pytest.mark.parametrize("dag,W,E,D,steps,pipe,tail,graph", TAIL_CASES)
def test_tail_grid_matches_oracle(oracle, product, monkeypatch, dag, W, E, D, steps, pipe, tail, graph):
    """One CTA per world with at least two rounds of worlds per CTA slot: the last round runs as a second grid of
    one-world CTAs beside the persistent grid (two launches, one completion ticket).  Same bit-exact bar, and the
    statistics count every world once."""
    monkeypatch.setenv("FIMO_B200_TEAM", "1")
    if pipe:
        monkeypatch.setenv("FIMO_B200_PIPE", pipe)
    if tail:
        monkeypatch.setenv("FIMO_B200_TAIL", tail)
    kw = {} if dag == "chain3" else {"dag": dag}
    got, want, info = run_both(oracle, product, W, E, D, steps, synth.RigOptions(fused=True, graph=graph), **kw)
    assert info.fused_world_step == 1
    assert info.kernel_launches_per_generation == 2 and info.nodes == 1  # dag16: prev_copy is inside the fused kernel too
    problems = synth.compare_outputs(got, want, FLOAT_RTOL)
    assert not problems, "\n".join(problems)
    st = got["stats"].reshape(-1).view(np.uint64)
    assert st[abi.STAT_STEPS] == W * steps and st[abi.STAT_WORLDS] == W and st[abi.STAT_ENTITIES] == W * E


@pytest.mark.parametrize("dag,disable", [("dag8", True), ("chain3", False)])
def test_tail_grid_off(oracle, product, monkeypatch, dag, disable):
    """FIMO_B200_TAIL=0 keeps the 8-system set on one grid; the 3-system chain is on one grid by default."""
    monkeypatch.setenv("FIMO_B200_TEAM", "1")
    if disable:
        monkeypatch.setenv("FIMO_B200_TAIL", "0")
    got, want, info = run_both(oracle, product, 1300, 64, 4, 2, dag=dag)
    assert info.fused_world_step == 1 and info.kernel_launches_per_generation == 1
    problems = synth.compare_outputs(got, want, FLOAT_RTOL)
    assert not problems, "\n".join(problems)


def test_large_world_per_level_propagate(oracle, product):
    """configs[2] world shape (64k entities, depth 8), one kernel per system: the world is too large for one
    CTA's shared memory, so propagate runs as one launch per hierarchy level."""
    got, want, info = run_both(oracle, product, 2, 65536, 8, 3, synth.RigOptions(fused=False))
    assert info.fused_world_step == 0 and info.kernel_launches_per_generation == 1 + 8 + 1
    problems = synth.compare_outputs(got, want, FLOAT_RTOL)
    assert not problems, "\n".join(problems)


def test_large_world_fused_pipeline(oracle, product, monkeypatch):
    """Same shape through the fused pipelined kernel: parents beyond the shared-memory window travel through the 48-byte row buffer in L2."""
    monkeypatch.setenv("FIMO_B200_TEAM", "1")
    got, want, info = run_both(oracle, product, 3, 65536, 8, 3)
    assert info.fused_world_step == 1 and info.kernel_launches_per_generation == 1
    problems = synth.compare_outputs(got, want, FLOAT_RTOL)
    assert not problems, "\n".join(problems)


def test_full_size_properties(product):
    """BASELINE configs[1] at full size (4096 x 1k): size-independent properties instead of the oracle."""
    W, E, D, K = 4096, 1024, 4, 4
    rig = synth.WorldStepRig(product, W, E, D)
    try:
        rig.step(K)
        out = rig.download()
        host = rig.host
    finally:
        rig.close()
    counts = out["vis_count"][:, 0, 0].astype(np.int64)
    bits = np.unpackbits(out["vis_mask"][:, 0, :].view(np.uint8), axis=1, bitorder="little")
    assert np.array_equal(bits.sum(axis=1), counts)                       # count == popcount(mask)
    st = out["stats"].reshape(-1).view(np.uint64)
    assert st[abi.STAT_VISIBLE] == counts.sum() and st[abi.STAT_ENTITIES] == W * E
    assert st[abi.STAT_MASK_SUM] == out["vis_mask"].astype(np.uint64).sum()  # checksum of checksums
    assert st[abi.STAT_STEPS] == W * K
    for w in (0, 1, 2047, 4095):                                          # list == ascending indices of the mask
        idx = np.nonzero(bits[w])[0].astype(np.uint32)
        assert np.array_equal(out["vis_list"][w, 0, :counts[w]], idx)
    assert synth.integrate_matches(host["pos"], host["vel"], K, out["pos"])   # K x pos = fma(vel, dt, pos)
    roots = host["parent"][:, 0, :] == abi.PARENT_ROOT                    # roots: world translation == pos
    for k in range(3):
        assert np.array_equal(out["world"][:, 4 * k + 3, :][roots], out["pos"][:, k, :][roots])


def run_full_size(oracle, product, W, E, D, K, dag):
    """One BASELINE-size batch through the product with the DEFAULT launch plan (no FIMO_B200_* overrides), then the
    same generations on the oracle (stepped in place, in world slices on the host threads): outputs and executor info."""
    data = synth.generate(W, E, D, dag=dag)
    dev = synth.WorldStepRig(product, W, E, D, data=data, dag=dag)  # uploads `data`; never writes it
    try:
        info = dev.executor_info()
        dev.step(K)
        got = dev.download()
    finally:
        dev.close()
    want = synth.step_in_slices(oracle, data, E, D, K, dag=dag)
    return got, want, info


FULL_SIZE_CASES = [
    # (id, dag, worlds, entities, depth, steps, kernel launches per generation of the default plan)
    ("C2", "chain3", 4096, 1024, 4, 16, 1),    # BASELINE configs[1]: grid 586 x 7 worlds per CTA
    ("C3", "dag8", 256, 65536, 8, 4, 2),        # configs[2]: CTA teams per world + list compaction
    ("C4-per-gpu", "chain3", 4096, 4096, 4, 4, 1),  # configs[3] = 32768 x 4k over 8 GPUs: the shard one GPU steps
    ("C5-per-gpu", "dag16", 1024, 1024, 4, 16, 1),  # configs[4]: 16-system DAG, 1 M entities per GPU: ONE kernel for all 16 systems
    ("dag8-4096", "dag8", 4096, 1024, 4, 4, 2),  # persistent grid + tail grid at bench size
    ("dag16-4096", "dag16", 4096, 1024, 4, 3, 1),
]


@pytest.mark.parametrize("name,dag,W,E,D,K,launches", FULL_SIZE_CASES, ids=[c[0] for c in FULL_SIZE_CASES])
def test_baseline_configs_at_full_size_match_oracle(oracle, product, monkeypatch, name, dag, W, E, D, K, launches):
    """Every BASELINE.json config at its STATED size, on the launch plan bench.py times (the plan is a function of the
    batch: worlds per CTA, tail-grid split, team choice — world_step.cu: world_step_plan), bit for bit against the
    oracle: every resource any system writes, after K generations."""
    for var in ("FIMO_B200_TEAM", "FIMO_B200_PIPE", "FIMO_B200_TAIL", "FIMO_B200_PREFETCH", "FIMO_B200_NO_SET2", "FIMO_B200_GENS"):
        monkeypatch.delenv(var, raising=False)
    got, want, info = run_full_size(oracle, product, W, E, D, K, dag)
    assert info.fused_world_step == 1
    if launches is not None:
        assert info.kernel_launches_per_generation == launches
    problems = synth.compare_outputs(got, want, FLOAT_RTOL)
    assert not problems, "\n".join(problems)
    st = got["stats"].reshape(-1).view(np.uint64)
    assert st[abi.STAT_STEPS] == W * K and st[abi.STAT_WORLDS] == W and st[abi.STAT_ENTITIES] == W * E
    counts = got["vis_count"][:, 0, 0].astype(np.int64)
    assert st[abi.STAT_VISIBLE] == counts.sum()


QUEUED_CASES = [
    # (dag, worlds, entities, depth, queued generations, FIMO_B200_GENS or None)
    ("chain3", 1200, 2048, 4, 37, None),   # 3 worlds per CTA: generations queued behind two launches share launches (up to 16 each)
    ("chain3", 600, 2048, 4, 21, "3"),     # T = 256: 3 worlds x 8 chunks per CTA; at most 3 generations per launch
    ("dag8", 1300, 1280, 4, 19, None),     # 10 chunks per world; single launches use persistent + tail grid, shared launches one grid over all worlds (no tail grid)
    ("dag16", 700, 1024, 4, 13, None),     # all 16 systems in the kernel, prev_copy included (reads the previous generation's matrices)
    ("chain3", 592, 256, 4, 12, None),     # 2 chunks per CTA < ring + 2: the plan refuses multi-generation launches
]


@pytest.mark.parametrize("dag,W,E,D,K,gens", QUEUED_CASES)
def test_queued_generations_share_launches_and_match_oracle(oracle, product, monkeypatch, dag, W, E, D, K, gens):
    """K generations scheduled back to back without waiting: once two launches are on the stream the following plain
    generations leave together in ONE launch (the ring runs across generation boundaries, every generation publishes
    its own statistics).  Same bit-exact bar after K generations; every generation's statistics land in the ring; the
    generation counter advances by one per schedule and the last fence fires after all of them."""
    import torch

    monkeypatch.setenv("FIMO_B200_TEAM", "1")
    if gens:
        monkeypatch.setenv("FIMO_B200_GENS", gens)
    data = synth.generate(W, E, D, dag=dag)
    ref = synth.WorldStepRig(oracle, W, E, D, data={k: v.copy() for k, v in data.items()}, dag=dag)
    per_step = []
    try:
        for _ in range(K):
            ref.step(1)
            per_step.append(ref.stats().copy())
        want = ref.download()
    finally:
        ref.close()
    ring = torch.zeros((K, abi.STAT_WORDS), dtype=torch.int64, device="cuda")
    opts = synth.RigOptions(stats_ring=ring.data_ptr(), stats_ring_slots=K)
    rig = synth.WorldStepRig(product, W, E, D, data={k: v.copy() for k, v in data.items()}, options=opts, dag=dag)
    try:
        info = rig.executor_info()
        before = rig.kernels_launched()
        rig.schedule(K)
        rig.wait()
        launched = rig.kernels_launched() - before
        got = rig.download()
        snapshots = ring.cpu().numpy().astype(np.uint64)
    finally:
        rig.close()
    expect_multi = E >= 5 * 128 and W >= 592
    assert (info.max_generations_per_launch > 1) == expect_multi, info.max_generations_per_launch
    if gens:
        assert info.max_generations_per_launch == int(gens)
    if expect_multi:
        assert launched < info.kernel_launches_per_generation * K, (launched, K)  # some generations shared a launch
    else:
        assert launched == info.kernel_launches_per_generation * K
    problems = synth.compare_outputs(got, want, FLOAT_RTOL)
    assert not problems, "\n".join(problems)
    for g in range(K):
        assert np.array_equal(snapshots[g][:5], per_step[g][:5]), (g, snapshots[g], per_step[g])


def test_generations_pipeline_and_fences(product):
    """schedule is asynchronous; K queued generations run in order; the signal fence fires after the last."""
    rig = synth.WorldStepRig(product, 32, 256, 4)
    try:
        rig.schedule(25)
        rig.wait()
        assert rig.stats()[abi.STAT_STEPS] == 32 * 25
        # wait_on: a generation gated on an unsignaled fence must not run until it is signaled
        from fimo_b200.worlds import Fence
        gate, done = Fence(product), Fence(product)
        rig.group.schedule(wait_on=[gate], signal=done)
        import time
        time.sleep(0.05)
        assert not done.is_signaled()
        gate.signal()
        done.wait()
        assert rig.stats()[abi.STAT_STEPS] == 32 * 26
    finally:
        rig.close()


def test_host_upload_between_steps(oracle, product):
    """world_lock_resources + upload of a per-step input (new frustum) is honoured by the next step."""
    W, E, D = 8, 512, 4
    data = synth.generate(W, E, D)
    other = synth.generate(W, E, D, world0=1000)["frustum"]
    outs = []
    for lib in (oracle, product):
        rig = synth.WorldStepRig(lib, W, E, D, data={k: v.copy() for k, v in data.items()})
        try:
            rig.step(2)
            rig.upload("frustum", other)
            rig.step(1)
            outs.append(rig.download())
        finally:
            rig.close()
    problems = synth.compare_outputs(outs[1], outs[0], FLOAT_RTOL)
    assert not problems, "\n".join(problems)


@pytest.mark.parametrize("case_name", ["group_schedule_counts_to_five", "context_generation_increments", "resource_unique_lock"])
def test_reference_goldens_on_gpu_box(product, case_name):
    import reference_goldens as G

    getattr(G, case_name)(product)


def test_stream_ordered_transfers_match_locked_host_access(oracle, product):
    """fimo_b200_group_transfer: uploads / downloads queued between generations give the same result as the
    lock-copy-unlock protocol on the oracle, and do not consume generation numbers."""
    from fimo_b200.worlds import Fence

    W, E, D = 8, 512, 4
    data = synth.generate(W, E, D)
    other = synth.generate(W, E, D, world0=1000)["frustum"]
    ref = synth.WorldStepRig(oracle, W, E, D, data={k: v.copy() for k, v in data.items()})
    try:
        ref.step(2)
        ref.upload("frustum", other)
        ref.step(3)
        want = ref.download()
    finally:
        ref.close()
    rig = synth.WorldStepRig(product, W, E, D, data={k: v.copy() for k, v in data.items()})
    try:
        got = {name: np.empty_like(rig.host[name]) for name in synth.OUTPUTS}
        rig.schedule(2, signal_last=False)
        rig.upload_async("frustum", other)          # ordered after generation 2, before generation 3
        rig.schedule(3, signal_last=False)
        fence = Fence(product)
        for i, name in enumerate(synth.OUTPUTS):
            rig.download_async(name, got[name], fence if i == len(synth.OUTPUTS) - 1 else None)
        fence.wait()
        assert rig.stats()[abi.STAT_STEPS] == W * 5
    finally:
        rig.close()
    problems = synth.compare_outputs(got, want, FLOAT_RTOL)
    assert not problems, "\n".join(problems)


@pytest.mark.parametrize("pinned", [False, True, "misaligned"], ids=["pageable", "pinned", "pinned-misaligned"])
def test_schedule_io_steps_in_flight_match_oracle(oracle, product, pinned):
    """fimo_b200_group_schedule_io: generations with their own input uploads and result downloads, several in
    flight, each with its own host buffers: every step's downloaded result equals the oracle's after the same
    uploads, and generation numbers advance by one per call.  pinned: page-locked buffers (fimo_b200_host_alloc) —
    the whole batch of small copies is one kernel that addresses host memory directly, uploads staged through HBM
    slots on the executor's copy stream while earlier generations run; misaligned: the same with host pointers that
    are only 4-byte aligned (32-bit copy path); pageable: the copy engine."""
    from fimo_b200.worlds import Fence

    W, E, D, K = 8, 512, 4, 9
    data = synth.generate(W, E, D)
    frusta = [synth.generate(W, E, D, world0=100 * (i + 1))["frustum"] for i in range(K)]
    dts = [(data["dt"] * np.float32(1.0 + 0.25 * i)).astype(np.float32) for i in range(K)]
    locked = []

    def buffer_like(a):
        if not pinned:
            return np.empty_like(a)
        skew = 4 if pinned == "misaligned" else 0
        ptr = product.fimo_b200_host_alloc(a.nbytes + skew)
        assert ptr
        locked.append(ptr)
        import ctypes as C
        return np.ctypeslib.as_array(C.cast(ptr + skew, C.POINTER(C.c_uint32)), shape=(a.size,)).view(a.dtype).reshape(a.shape)

    ref = synth.WorldStepRig(oracle, W, E, D, data={k: v.copy() for k, v in data.items()})
    want = []
    try:
        for i in range(K):
            ref.upload("frustum", frusta[i])
            ref.upload("dt", dts[i])
            ref.step(1)
            want.append(ref.download(("vis_count", "stats", "vis_mask")))
    finally:
        ref.close()
    rig = synth.WorldStepRig(product, W, E, D, data={k: v.copy() for k, v in data.items()})
    try:
        gen0 = rig.group.generation() if hasattr(rig.group, "generation") else None
        fences = [Fence(product) for _ in range(K)]
        got = [{n: buffer_like(rig.host[n]) for n in ("vis_count", "stats", "vis_mask")} for _ in range(K)]
        lists = []
        for i in range(K):  # all K steps queued before any is waited for
            src, src_dt = buffer_like(frusta[i]), buffer_like(dts[i])
            src[...], src_dt[...] = frusta[i], dts[i]
            up, down = rig.io_list({"frustum": src, "dt": src_dt}), rig.io_list(got[i])
            lists.append((up, down))
            rig.schedule_io(up, down, fences[i])
        for f in fences:
            f.wait()
        for i in range(K):
            problems = synth.compare_outputs(got[i], want[i], 0.0)
            assert not problems, f"step {i}:\n" + "\n".join(problems)
        assert rig.stats()[abi.STAT_STEPS] == W * K
        if gen0 is not None:
            assert rig.group.generation() == gen0 + K
        got = [{n: v.copy() for n, v in g.items()} for g in got]
    finally:
        rig.close()
        for ptr in locked:
            product.fimo_b200_host_free(ptr)


def test_schedule_io_result_fed_back_through_host_memory(oracle, product):
    """A step's download and the next step's upload share ONE pinned host buffer, all steps queued without waiting:
    positions read back by step i become the velocities uploaded for step i+1.  The copies must stay in stream
    order (the staged, early PCIe read of an upload would see the buffer before the earlier download has landed)."""
    import ctypes as C
    from fimo_b200.worlds import Fence

    W, E, D, K = 6, 768, 4, 7
    data = synth.generate(W, E, D)
    ref = synth.WorldStepRig(oracle, W, E, D, data={k: v.copy() for k, v in data.items()})
    try:
        x = data["vel"].copy()
        for _ in range(K):
            ref.upload("vel", x)
            ref.step(1)
            x = ref.download(("pos",))["pos"]
        want = ref.download()
    finally:
        ref.close()
    ptr = product.fimo_b200_host_alloc(data["vel"].nbytes)
    assert ptr
    rig = synth.WorldStepRig(product, W, E, D, data={k: v.copy() for k, v in data.items()})
    try:
        buf = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint32)), shape=(data["vel"].size,)).view(np.float32).reshape(data["vel"].shape)
        buf[...] = data["vel"]
        fences = [Fence(product) for _ in range(K)]
        up, down = rig.io_list({"vel": buf}), rig.io_list({"pos": buf})
        for i in range(K):
            rig.schedule_io(up, down, fences[i])
        for f in fences:
            f.wait()
        got = rig.download()
        assert np.array_equal(buf.view(np.uint32), want["pos"].view(np.uint32))
    finally:
        rig.close()
        product.fimo_b200_host_free(ptr)
    problems = synth.compare_outputs(got, want, 0.0)
    assert not problems, "\n".join(problems)


@pytest.mark.parametrize("pack", ["beside-next-generation", "in-stream"])
@pytest.mark.parametrize("W,E,D,K,margin", [(24, 1024, 4, 7, 1.0), (700, 96, 4, 5, 1.0), (3, 8192, 8, 4, 1.0), (24, 1024, 4, 3, 0.5)])
def test_readback_returns_what_cull_produces(oracle, product, monkeypatch, W, E, D, K, margin, pack):
    """fimo_b200_group_schedule_readback: every step uploads its frusta and gets back, in ONE packed pinned buffer, the
    per-world visible counts, their exclusive offsets, the masks and the ordered visible lists back to back — all K
    steps queued before the first wait, each with its own buffer; every buffer equals the oracle's outputs of that
    step.  margin < 1: the list does not fit -> `truncated`, the prefix that fits is still exact.
    pack: one-CTA-per-world plans leave a second copy of the cull outputs and pack it on the read-back stream beside the
    next generation (default); "in-stream" packs from the resources in stream order (what CTA-team plans always do).  Either
    way the resources themselves hold the last step's outputs afterwards."""
    import ctypes as C
    from fimo_b200.worlds import Fence

    if pack == "in-stream":
        monkeypatch.setenv("FIMO_B200_READBACK_IN_STREAM", "1")
    else:
        monkeypatch.delenv("FIMO_B200_READBACK_IN_STREAM", raising=False)

    data = synth.generate(W, E, D)
    frusta = [synth.generate(W, E, D, world0=100 * (i + 1))["frustum"] for i in range(K)]
    ref = synth.WorldStepRig(oracle, W, E, D, data={k: v.copy() for k, v in data.items()})
    want = []
    try:
        for i in range(K):
            ref.upload("frustum", frusta[i])
            ref.step(1)
            want.append(ref.download(("vis_count", "vis_mask", "vis_list")))
    finally:
        ref.close()
    rig = synth.WorldStepRig(product, W, E, D, data={k: v.copy() for k, v in data.items()})
    locked = []
    try:
        capacity = int(margin * max(int(w["vis_count"].sum()) for w in want))
        words = rig.readback_words(capacity)

        def pinned(n_words):
            ptr = product.fimo_b200_host_alloc(4 * n_words)
            assert ptr
            locked.append(ptr)
            return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint32)), shape=(n_words,))

        bufs, fences, keep = [pinned(words) for _ in range(K)], [Fence(product) for _ in range(K)], []
        for i in range(K):
            src = pinned(frusta[i].size).view(np.float32).reshape(frusta[i].shape)
            src[...] = frusta[i]
            up, desc = rig.io_list({"frustum": src}), rig.readback_desc(bufs[i])
            keep.append((up, desc))
            rig.schedule_readback(up, desc, fences[i])
        for f in fences:
            f.wait()
        for i in range(K):
            rb = rig.parse_readback(bufs[i])
            counts = want[i]["vis_count"][:, 0, 0]
            assert rb["n_worlds"] == W and rb["entities"] == E and rb["generation"] == i and rb["total"] == int(counts.sum())
            assert np.array_equal(rb["counts"], counts) and np.array_equal(rb["offsets"], np.cumsum(counts) - counts)
            assert np.array_equal(rb["mask"], want[i]["vis_mask"][:, 0, :])
            full = np.concatenate([want[i]["vis_list"][w, 0, :counts[w]] for w in range(W)])
            assert rb["truncated"] == (1 if rb["total"] > capacity else 0)
            assert np.array_equal(rb["list"], full[:capacity]), i
        last = rig.download(("vis_count", "vis_mask", "vis_list"))  # the resources, not the shadow copy
        counts = want[-1]["vis_count"][:, 0, 0]
        assert np.array_equal(last["vis_count"], want[-1]["vis_count"]) and np.array_equal(last["vis_mask"], want[-1]["vis_mask"])
        for w in range(W):
            assert np.array_equal(last["vis_list"][w, 0, :counts[w]], want[-1]["vis_list"][w, 0, :counts[w]])
    finally:
        rig.close()
        for ptr in locked:
            product.fimo_b200_host_free(ptr)


@pytest.mark.parametrize("mode", ["fused", "graph", "streams"])
@pytest.mark.parametrize("W,E,D,steps", [(16, 1024, 4, 6), (40, 260, 5, 4), (3, 8192, 8, 3)])
def test_mixed_host_and_device_systems_match_oracle(oracle, product, W, E, D, steps, mode):
    """host systems (plain `system_run` callbacks on the pool) and device systems in ONE group.
    `host_camera` rewrites every world's frustum before `cull` (ordered by the W/R resource conflict only) and
    `host_census` sums the per-world visible counts after `cull` into a host resource.  On the product the host
    systems run from host nodes of the generation's CUDA graph and see pinned mirrors of the device resources;
    every output, the rewritten frusta and the census equal the CPU oracle's."""
    data = synth.generate(W, E, D)
    names = synth.OUTPUTS + ("frustum",)
    ref = synth.WorldStepRig(oracle, W, E, D, data={k: v.copy() for k, v in data.items()}, host_systems=True)
    try:
        ref.step(steps)
        want, want_census = ref.download(names), ref.census()
        assert not ref.host_errors, ref.host_errors
    finally:
        ref.close()
    opts = synth.RigOptions(fused=(mode == "fused"), graph=(mode != "streams"))
    dev = synth.WorldStepRig(product, W, E, D, data={k: v.copy() for k, v in data.items()}, options=opts, host_systems=True)
    try:
        info = dev.executor_info()
        dev.step(steps)
        got, got_census = dev.download(names), dev.census()
        assert not dev.host_errors, dev.host_errors
    finally:
        dev.close()
    assert info.systems == 5 and info.device_systems == 3
    assert info.nodes == (3 if mode == "fused" and E % 4 == 0 else 5)  # camera -> fused world-step -> census
    assert info.fused_world_step == (1 if mode == "fused" else 0)
    problems = synth.compare_outputs(got, want, FLOAT_RTOL)
    assert not problems, "\n".join(problems)
    assert got_census == want_census and got_census[1] == steps
    assert got_census[2] == int(got["vis_count"].sum())


@pytest.mark.parametrize("dag,W,E,D,fused", [("chain3", 24, 512, 4, True), ("chain3", 24, 512, 4, False), ("chain3", 3, 8192, 8, True),
                                            ("dag16", 700, 128, 4, True), ("dag8", 9, 1000, 4, True),
                                            ("dag8", 1300, 64, 4, True)])  # persistent + tail grid share one ticket
def test_stats_ring_holds_every_generation(product, dag, W, E, D, fused):
    """FimoB200_GroupDeviceExt.stats_ring: the kernel that publishes the statistics also stores them in slot
    generation % slots — K generations queued back to back leave the same words a step-by-step run reads from `stats`."""
    import torch

    slots, K = 4, 7
    data = synth.generate(W, E, D, dag=dag)
    rig = synth.WorldStepRig(product, W, E, D, data={k: v.copy() for k, v in data.items()}, options=synth.RigOptions(fused=fused), dag=dag)
    try:
        want = []
        for _ in range(K):
            rig.step(1)
            want.append(rig.stats().copy())
    finally:
        rig.close()
    ring = torch.zeros((slots, abi.STAT_WORDS), dtype=torch.int64, device="cuda")
    opts = synth.RigOptions(fused=fused, stats_ring=ring.data_ptr(), stats_ring_slots=slots)
    rig = synth.WorldStepRig(product, W, E, D, data={k: v.copy() for k, v in data.items()}, options=opts, dag=dag)
    try:
        rig.step(K)  # all K queued, then one wait
        got = ring.cpu().numpy().astype(np.uint64)
    finally:
        rig.close()
    for g in range(K - slots, K):
        assert np.array_equal(got[g % slots], want[g]), (g, got[g % slots], want[g])


def test_trace_sink_reports_every_device_generation(product):
    """with a sink installed the completion thread reports each finished device generation (label,
    generation number, device time from the bracketing CUDA events, host timestamps, kernel count)."""
    import ctypes as C

    events = []

    def sink(_user, ev):
        e = ev.contents
        events.append((C.string_at(e.group_label, e.group_label_len).decode(), int(e.generation), float(e.device_ms),
                       int(e.host_enqueue_ns), int(e.host_complete_ns), int(e.kernels), int(e.fused_world_step)))

    cb = abi.TraceSinkFn(sink)
    rig = synth.WorldStepRig(product, 64, 1024, 4)
    try:
        rig.step(2)  # before the sink: not reported
        product.fimo_worlds_b200_set_trace_sink(cb, None)
        rig.step(5)
    finally:
        product.fimo_worlds_b200_set_trace_sink(abi.TraceSinkFn(), None)
        rig.close()
    assert [e[1] for e in events] == [2, 3, 4, 5, 6]
    assert all(e[0] == "world-step" and e[5] in (1, 2) and e[6] == 1 for e in events)  # 2 = team kernel + list compaction
    assert all(0.0 < e[2] < 50.0 and 0 < e[3] <= e[4] for e in events)


WIND_LIB = __import__("os").path.join(__import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))),
                                      "examples", "wind_system", "libwind_system.so")


@pytest.mark.parametrize("placement,mode", [("between", "fused"), ("between", "graph"), ("between", "streams"),
                                            ("front", "fused"), ("front", "graph"), ("front", "streams")])
@pytest.mark.parametrize("W,E,D,steps", [(24, 1024, 4, 5), (700, 260, 5, 3)])
def test_user_device_system_through_device_entry(oracle, product, placement, mode, W, E, D, steps):
    """A system the library has never heard of (examples/wind_system, its own .so; builtin_id 0) registered through the
    unchanged SystemDescriptor + the device extension on `next` (systems.h:51-114): its `device_entry` launches its
    kernel on the generation's stream — captured into the CUDA graph, or launched live — ordered against the built-ins by
    `before` edges and resource conflicts only.  "between": after `forces`, before `integrate_lin` — the fused set is no
    longer convex, so every system runs as its own node; "front": in front of `forces` — the 8 built-ins still fuse
    AROUND it (2 nodes).  Same math as its host `system_run` on the oracle: every output bit-exact."""
    data = synth.generate(W, E, D, dag="dag8")
    ref = synth.WorldStepRig(oracle, W, E, D, data={k: v.copy() for k, v in data.items()}, dag="dag8", wind=(WIND_LIB, placement))
    try:
        ref.step(steps)
        want = ref.download()
        assert ref.wind_counters() == (0, steps)  # the CPU path ran system_run
    finally:
        ref.close()
    opts = synth.RigOptions(fused=(mode == "fused"), graph=(mode != "streams"))
    dev = synth.WorldStepRig(product, W, E, D, data={k: v.copy() for k, v in data.items()}, options=opts, dag="dag8", wind=(WIND_LIB, placement))
    try:
        info = dev.executor_info()
        dev.step(steps)
        got = dev.download()
        launches, host_runs = dev.wind_counters()
    finally:
        dev.close()
    assert info.systems == 9 and info.device_systems == 9 and host_runs == 0  # its system_run is never called on the product
    if mode == "fused" and placement == "front" and E % 4 == 0:
        assert info.fused_world_step == 1 and info.nodes == 2
    else:
        assert info.fused_world_step == 0 and info.nodes == 9
    assert launches == (steps if mode == "streams" else 0)  # live launches per generation; under capture the graph replays the kernel
    problems = synth.compare_outputs(got, want, FLOAT_RTOL)
    assert not problems, "\n".join(problems)


DAG_CASES = [
    # (dag, worlds, entities, depth, steps)
    ("dag8", 1, 1024, 4, 8),
    ("dag8", 3, 65536, 8, 2),     # BASELINE configs[2] world shape
    ("dag8", 37, 1000, 8, 4),     # ragged
    ("dag16", 4, 4096, 4, 4),
    ("dag16", 61, 260, 5, 6),
    ("dag16", 2, 33, 3, 3),
    ("dag8", 700, 128, 4, 3),     # more worlds than CTA slots, one chunk per world
    ("dag16", 5, 8192, 8, 2),
    ("dag16", 1200, 256, 4, 3),   # more worlds than CTA slots: 15 systems in the world-step kernel, several worlds per CTA
    ("dag16", 300, 1024, 4, 3),   # BASELINE configs[4] world shape
]


@pytest.mark.parametrize("dag,W,E,D,steps", DAG_CASES)
@pytest.mark.parametrize("mode", ["fused", "fused-one-cta-per-world", "graph", "streams"])
def test_system_dags_match_oracle(oracle, product, monkeypatch, dag, W, E, D, steps, mode):
    """The 8-system fan-out/fan-in DAG and the 16-system DAG with conflict-only orderings: every resource any
    system writes is bit-exact.  fused: the 8-system core (plus vel_clamp) runs inside the world-step kernel and
    the other systems stay graph nodes around it; graph / streams: one kernel per system."""
    if mode == "fused-one-cta-per-world":
        monkeypatch.setenv("FIMO_B200_TEAM", "1")
    fused = mode.startswith("fused")
    got, want, info = run_both(oracle, product, W, E, D, steps, synth.RigOptions(fused=fused, graph=(mode != "streams")), dag=dag)
    n = 8 if dag == "dag8" else 16
    assert info.systems == n and info.device_systems == n
    if fused and E % 4 == 0:
        # dag16: all 16 systems inside the world-step kernel (1 node); with CTA teams (few, large worlds) the 8-system
        # core + vel_clamp is fused and seven systems stay graph nodes (8 nodes)
        assert info.fused_world_step == 1 and info.nodes in ((1,) if dag == "dag8" else (1, 8))
        if dag == "dag16" and (mode == "fused-one-cta-per-world" or W >= 300):
            assert info.nodes == 1
    else:
        assert info.fused_world_step == 0 and info.nodes == n and info.edges >= n - 1
    assert info.uses_cuda_graph == (0 if mode == "streams" else 1)
    problems = synth.compare_outputs(got, want, FLOAT_RTOL)
    assert not problems, "\n".join(problems)
    st = got["stats"].reshape(-1).view(np.uint64)
    assert st[abi.STAT_STEPS] == W * steps
