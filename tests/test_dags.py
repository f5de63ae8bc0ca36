"""The 8-system fan-out / fan-in DAG (BASELINE.json configs[2]) and the 16-system DAG with mixed read/write
conflicts (configs[4]) on the CPU oracle: the scheduler's ordering contract (SystemGroup.zig:160-338 — rank by
explicit dependencies, then add order; conflicts W/W, R/W, W/R serialise in that order) checked against an
independent numpy evaluation of each system."""
import numpy as np
import pytest

from fimo_b200 import abi, synth

F32 = np.float32


def run(lib, dag, W, E, D, steps):
    rig = synth.WorldStepRig(lib, W, E, D, dag=dag)
    try:
        init = {k: v.copy() for k, v in rig.host.items()}
        rig.step(steps)
        return init, rig.download()
    finally:
        rig.close()


def fma(a, b, c):
    """fmaf of binary32 operands: the product is exact in binary64; the sum may double-round in the last bit, so
    callers compare with a 1-ulp allowance where an fma feeds the value."""
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(F32)


def close_ulp(a, b, ulps=1):
    a, b = a.astype(F32), b.astype(F32)
    return bool((np.abs(a.view(np.int32).astype(np.int64) - b.view(np.int32).astype(np.int64)) <= ulps).all())


@pytest.mark.parametrize("dag", ["dag8", "dag16"])
def test_fan_in_statistics_are_consistent(oracle, dag):
    W, E = 5, 300
    _, out = run(oracle, dag, W, E, 4, 3)
    gs = out["gstats"].reshape(-1)
    assert gs[abi.GSTAT_VISIBLE_LO:abi.GSTAT_VISIBLE_LO + 2].view(np.uint64)[0] == out["vis_count"].sum()  # cull -> stats
    assert gs[abi.GSTAT_WORLDS_LO:abi.GSTAT_WORLDS_LO + 2].view(np.uint64)[0] == W
    wa = out["waabb"]
    lo, hi = (wa[:, :3, :] - wa[:, 3:, :]).min(axis=2), (wa[:, :3, :] + wa[:, 3:, :]).max(axis=2)
    assert np.array_equal(out["wbounds"][:, :3, 0], lo) and np.array_equal(out["wbounds"][:, 3:, 0], hi)  # world_aabb -> bounds
    gb = gs[abi.GSTAT_BOUNDS:abi.GSTAT_BOUNDS + 6].view(F32)
    assert np.array_equal(gb[:3], lo.min(axis=0)) and np.array_equal(gb[3:], hi.max(axis=0))               # bounds -> stats
    bits = np.unpackbits(out["vis_mask"][:, 0, :].view(np.uint8), axis=1, bitorder="little")[:, :E]
    assert np.array_equal(bits.sum(axis=1), out["vis_count"][:, 0, 0])
    for w in range(W):
        assert np.array_equal(out["vis_list"][w, 0, :out["vis_count"][w, 0, 0]], np.nonzero(bits[w])[0])
    assert np.abs((out["rot"].astype(np.float64) ** 2).sum(axis=1) - 1.0).max() < 1e-6                    # integrate_ang renormalises


def test_dag8_order_forces_before_integrate(oracle):
    """pos after one step = fma(vel', dt, pos) with vel' = fma(g, dt, vel): forces ran before integrate_lin."""
    init, out = run(oracle, "dag8", 2, 128, 4, 1)
    dt, g = init["dt"], init["gravity"]
    vel1 = fma(np.broadcast_to(g, init["vel"].shape), np.broadcast_to(dt, init["vel"].shape), init["vel"])
    assert close_ulp(out["vel"], vel1)
    assert close_ulp(out["pos"], fma(out["vel"], np.broadcast_to(dt, out["vel"].shape), init["pos"]))


def test_dag16_conflict_only_orderings(oracle):
    """Orderings that no `before` edge states, only resource conflicts in rank-then-add order."""
    K = 3
    init, out = run(oracle, "dag16", 3, 160, 4, K)
    _, prev = run(oracle, "dag16", 3, 160, 4, K - 1)
    # prev_copy (S world) runs before propagate (X world): it holds the previous step's transforms
    assert np.array_equal(out["prev_world"].view(np.uint32), prev["world"].view(np.uint32))
    # forces -> vel_clamp (W/W on vel): the result is clamped, and never exceeds vmax
    vmax = init["vmax"]
    assert (np.abs(out["vel"]) <= vmax).all()
    v = init["vel"]
    for _ in range(K):
        v = fma(np.broadcast_to(init["gravity"], v.shape), np.broadcast_to(init["dt"], v.shape), v)
        v = np.minimum(np.maximum(v, -vmax), vmax)
    assert close_ulp(out["vel"], v, ulps=K)
    # sleep_flags reads the clamped velocity of this step
    a = (np.abs(out["vel"][:, 0, :]) + np.abs(out["vel"][:, 1, :])).astype(F32) + np.abs(out["vel"][:, 2, :])
    bits = np.unpackbits(out["sleep"][:, 0, :].view(np.uint8), axis=1, bitorder="little")[:, :160]
    assert np.array_equal(bits.astype(bool), a.astype(F32) < init["sleep_eps"][:, 0, :])
    # angvel_damp -> integrate_ang (W then R on angvel): K dampings
    w = init["angvel"]
    for _ in range(K):
        w = (w * init["damp"]).astype(F32)
    assert np.array_equal(out["angvel"].view(np.uint32), w.view(np.uint32))
    # lod follows world_aabb of this step; vis_bounds follows cull
    d = out["waabb"][:, :3, :] - init["campos"]
    d2 = fma(d[:, 2], d[:, 2], fma(d[:, 1], d[:, 1], (d[:, 0] * d[:, 0]).astype(F32)))
    lod = sum((d2 >= init["lod_dist2"][:, i, :]).astype(np.uint32) for i in range(3))
    assert (out["lod"][:, 0, :] != lod).mean() < 0.01  # double rounding in the numpy fma may flip a threshold case
    vb = np.unpackbits(out["vis_mask"][:, 0, :].view(np.uint8), axis=1, bitorder="little")[:, :160].astype(bool)
    for wi in range(3):
        if vb[wi].any():
            wa = out["waabb"][wi][:, vb[wi]]
            assert np.array_equal(out["vbounds"][wi, :3, 0], (wa[:3] - wa[3:]).min(axis=1))
            assert np.array_equal(out["vbounds"][wi, 3:, 0], (wa[:3] + wa[3:]).max(axis=1))


@pytest.mark.parametrize("dag", ["dag8", "dag16"])
def test_dag_is_deterministic(oracle, dag):
    _, a = run(oracle, dag, 4, 96, 8, 4)
    _, b = run(oracle, dag, 4, 96, 8, 4)
    assert not synth.compare_outputs(a, b, 0.0)
