"""The synthetic world generator and the oracle's systems on it (CPU)."""
import numpy as np

from fimo_b200 import abi, synth


def test_generator_is_shard_independent():
    whole = synth.generate(6, 96, 4, world0=0)
    part = synth.generate(2, 96, 4, world0=3)
    for name in ("pos", "vel", "rot", "scale", "parent", "aabb_c", "aabb_h", "frustum", "level_start"):
        assert np.array_equal(whole[name][3:5], part[name]), name


def test_hierarchy_is_level_sorted():
    for E, D in ((1024, 4), (4096, 4), (65536, 8), (100, 8), (5, 4), (3, 4)):
        d = synth.generate(2, E, D)
        ls = d["level_start"][0, :, 0]
        assert ls[0] == 0 and ls[-1] == E and (np.diff(ls.astype(np.int64)) >= 0).all()
        parent = d["parent"][:, 0, :]
        for l in range(D):
            lo, hi = int(ls[l]), int(ls[l + 1])
            p = parent[:, lo:hi]
            if l == 0 or ls[l] == ls[l - 1]:
                assert (p == abi.PARENT_ROOT).all() or l > 0
            else:
                assert ((p >= ls[l - 1]) & (p < ls[l])).all()


def test_quaternions_are_unit_and_planes_normalised():
    d = synth.generate(4, 512, 4)
    n = np.sqrt((d["rot"].astype(np.float64) ** 2).sum(axis=1))
    assert np.allclose(n, 1.0, atol=1e-6)
    fr = d["frustum"].reshape(4, 6, 4)
    assert np.allclose(np.linalg.norm(fr[:, :, :3], axis=2), 1.0, atol=1e-6)


def test_oracle_chain_properties(oracle):
    """Size-independent properties of the CPU definition: list is the ascending index list of the mask,
    count == popcount(mask), stats are the sums."""
    rig = synth.WorldStepRig(oracle, 5, 300, 4)
    try:
        rig.step(4)
        out = rig.download()
    finally:
        rig.close()
    total = 0
    mask_sum = 0
    for w in range(5):
        bits = np.unpackbits(out["vis_mask"][w, 0].view(np.uint8), bitorder="little")[:300]
        idx = np.nonzero(bits)[0]
        c = int(out["vis_count"][w, 0, 0])
        assert c == len(idx)
        assert np.array_equal(out["vis_list"][w, 0, :c], idx.astype(np.uint32))
        assert (out["vis_list"][w, 0, c:] == 0).all()
        total += c
        mask_sum += int(out["vis_mask"][w, 0].astype(np.uint64).sum())
    st = out["stats"].reshape(-1).view(np.uint64)
    assert st[abi.STAT_VISIBLE] == total and st[abi.STAT_ENTITIES] == 5 * 300
    assert st[abi.STAT_MASK_SUM] == mask_sum and st[abi.STAT_WORLDS] == 5 and st[abi.STAT_STEPS] == 20
    assert 0 < total < 5 * 300
    # integrate is K fused multiply-adds pos = fma(vel, dt, pos)
    d0 = synth.generate(5, 300, 4)
    assert synth.integrate_matches(d0["pos"], d0["vel"], 4, out["pos"])


def test_host_systems_in_the_chain_on_the_oracle(oracle):
    """The mixed-group rig on the CPU oracle: host_camera rolls the frusta once per generation, host_census runs
    after cull and sums exactly the counts cull just wrote."""
    W, E, D, K = 6, 256, 4, 3
    data = synth.generate(W, E, D)
    rig = synth.WorldStepRig(oracle, W, E, D, data={k: v.copy() for k, v in data.items()}, host_systems=True)
    try:
        rig.step(K)
        out = rig.download(("frustum", "vis_count"))
        visible, generations, last = rig.census()
        assert not rig.host_errors, rig.host_errors
    finally:
        rig.close()
    assert generations == K and last == int(out["vis_count"].sum()) and visible >= last
    assert np.array_equal(out["frustum"], np.roll(data["frustum"], K, axis=0))
