"""The oracle's ARITHMETIC, checked against an independent model (CPU only).

The reference has no transform / AABB / cull code, so GPU parity for the synthetic systems means
"CUDA == oracle/systems_ref.c" — two files written op for op by the same hand.  This test is the third opinion: a float64
NumPy model written from the DEFINITIONS (SURVEY.md §8d "Systems (3-chain)" / "Proposed 8-system DAG", the one-line
contracts in oracle/systems_ref.h), never from the C code: textbook quaternion -> rotation matrix, affine product with
matrix algebra, c' = W c + t, h' = |W| h, the six-plane test.

Every generation is checked on its own: the model is fed the oracle's state BEFORE the step (float32 promoted to
float64) and must reproduce the oracle's state AFTER it — floats within 1e-5 (relative to the magnitude of the data),
integers exactly except where the float64 model itself sits within 1e-3 of a decision threshold (counted, < 0.1 %).
A sign error, a transposed matrix or a swapped plane component in systems_ref.c fails here by orders of magnitude."""
import numpy as np
import pytest

from fimo_b200 import abi, synth

ROOT = 0xFFFFFFFF
RTOL = 1e-5
EMPTY = np.float32(3.402823466e+38)


def f64(a):
    return a.astype(np.float64)


def close(got, want, scale=None):
    """|got - want| <= RTOL * max(|want|, scale): relative to the data's magnitude (positions reach a few hundred)."""
    want = np.asarray(want, dtype=np.float64)
    scale = np.abs(want).max() if scale is None else scale
    return np.abs(f64(np.asarray(got)) - want) <= RTOL * np.maximum(np.abs(want), scale)


def rotation(q):
    """Textbook rotation matrix of a unit quaternion (x, y, z, w); q: [W, 4, E] -> [W, E, 3, 3]."""
    x, y, z, w = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    return np.stack([np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)], -1),
                     np.stack([2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)], -1),
                     np.stack([2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], -1)], -2)


def propagate(pos, rot, scale, parent, level_start):
    """world = root ? T R S : world[parent] o (T R S), levels in order.  Returns A [W, E, 3, 3], t [W, E, 3]."""
    W, _, E = pos.shape
    A = rotation(rot) * scale[:, 0, :, None, None]
    t = np.moveaxis(pos, 1, 2).copy()
    for w in range(W):
        ls = level_start[w, :, 0]
        for lvl in range(len(ls) - 1):
            idx = np.arange(ls[lvl], ls[lvl + 1])
            idx = idx[parent[w, 0, idx] != ROOT]
            p = parent[w, 0, idx]
            assert (p < idx).all()
            t[w, idx] = np.einsum("eij,ej->ei", A[w, p], t[w, idx]) + t[w, p]
            A[w, idx] = A[w, p] @ A[w, idx]
    return A, t


def matrices_of(world):
    """[W, 12, E] row-major 3x4 -> A [W, E, 3, 3], t [W, E, 3]."""
    m = np.moveaxis(f64(world), 1, 2).reshape(world.shape[0], world.shape[2], 3, 4)
    return m[..., :3], m[..., 3]


def slack(c, h, frustum):
    """Smallest plane slack n.c + d + |n|.h over the six planes; c, h: [W, E, 3]; frustum [W, 24, 1]."""
    pl = f64(frustum)[:, :, 0].reshape(-1, 6, 4)
    sd = np.einsum("wpk,wek->wep", pl[..., :3], c) + pl[:, None, :, 3] + np.einsum("wpk,wek->wep", np.abs(pl[..., :3]), h)
    return sd.min(-1)


def bits(mask, E):
    return np.unpackbits(mask[:, 0, :].view(np.uint8), axis=1, bitorder="little")[:, :E].astype(bool)


def decisions_match(got_bool, model_margin, what, tally):
    """A yes/no the oracle took in float32 against the model's float64 margin (>= 0 means yes).  Away from the threshold
    they must agree; entities the model itself places within 1e-3 of it may go either way and are counted."""
    near = np.abs(model_margin) < 1e-3
    bad = (got_bool != (model_margin >= 0)) & ~near
    assert not bad.any(), f"{what}: {int(bad.sum())} decisions differ away from the threshold"
    tally[0] += int(near.sum())
    tally[1] += got_bool.size


def check_cull(after, names, c, h, frustum, E, tally):
    mask, lst, cnt = (after[n] for n in names)
    vis = bits(mask, E)
    decisions_match(vis, slack(c, h, frustum), names[0], tally)
    assert np.array_equal(cnt[:, 0, 0], vis.sum(1))
    for w in range(vis.shape[0]):  # the list is the ascending indices of the set bits
        assert np.array_equal(lst[w, 0, :cnt[w, 0, 0]], np.nonzero(vis[w])[0])
    if E % 32:
        assert not (mask[:, 0, -1] >> np.uint32(E % 32)).any()  # no bits past the last entity
    return vis


def bounds_of(c, h, select=None):
    lo, hi = c - h, c + h
    if select is not None:
        lo, hi = np.where(select[..., None], lo, np.inf), np.where(select[..., None], hi, -np.inf)
    return np.concatenate([lo.min(1), hi.max(1)], axis=1)  # [W, 6]


def check_bounds(got, want):
    got, empty = f64(got[:, :, 0]), ~np.isfinite(want)
    assert np.array_equal(np.abs(got[empty]), np.full(int(empty.sum()), float(EMPTY)))  # nothing selected: +-FLT_MAX
    assert close(got[~empty], want[~empty]).all()


def check_step(dag, before, after, E, tally):
    dt = f64(before["dt"])  # [W, 1, 1]
    vel, rot = f64(before["vel"]), f64(before["rot"])
    if dag != "chain3":  # forces: vel += g dt
        vel = vel + f64(before["gravity"]) * dt
        if dag == "dag16":  # vel_clamp between forces and integrate_lin
            vmax = f64(before["vmax"])
            vel = np.clip(vel, -vmax, vmax)
        assert close(after["vel"], vel).all(), "forces / vel_clamp"
        om = f64(before["angvel"])
        if dag == "dag16":  # angvel_damp before integrate_ang
            om = om * f64(before["damp"])
            assert close(after["angvel"], om).all(), "angvel_damp"
        # integrate_ang: q += (dt / 2) (omega, 0) (x) q, then normalise.  Quaternion product, vector part first:
        # (a, 0)(v, s) = (a s + a x v, -a . v)
        v, s = rot[:, :3], rot[:, 3:4]
        dq = np.concatenate([om * s + np.cross(om, v, axis=1), -(om * v).sum(1, keepdims=True)], axis=1)
        rot = rot + 0.5 * dt * dq
        rot = rot / np.sqrt((rot * rot).sum(1, keepdims=True))
        assert close(after["rot"], rot, scale=1.0).all(), "integrate_ang"
        assert np.allclose((f64(after["rot"]) ** 2).sum(1), 1.0, atol=1e-6), "rot is not unit length"
    pos = f64(before["pos"]) + vel * dt  # integrate
    assert close(after["pos"], pos).all(), "integrate"

    A, t = propagate(pos, rot, f64(before["scale"]), before["parent"], before["level_start"])
    gA, gt = matrices_of(after["world"])
    assert close(gA, A, scale=1.0).all() and close(gt, t).all(), "propagate"
    # the 3x3 part of every world matrix is a rotation times a positive uniform scale: M M^T = s^2 I, det > 0
    gram = gA @ np.swapaxes(gA, -1, -2)
    s2 = np.trace(gram, axis1=-2, axis2=-1) / 3.0
    assert (np.abs(gram - s2[..., None, None] * np.eye(3)) <= 1e-5 * s2[..., None, None]).all(), "M M^T != s^2 I"
    assert (np.linalg.det(gA) > 0).all()

    c = np.einsum("weij,wej->wei", A, np.moveaxis(f64(before["aabb_c"]), 1, 2)) + t  # c' = W c + t
    h = np.einsum("weij,wej->wei", np.abs(A), np.moveaxis(f64(before["aabb_h"]), 1, 2))  # h' = |W| h
    vis = check_cull(after, ("vis_mask", "vis_list", "vis_count"), c, h, before["frustum"], E, tally)
    st = after["stats"].reshape(-1).view(np.uint64)
    n_worlds = pos.shape[0]
    assert st[abi.STAT_VISIBLE] == vis.sum() and st[abi.STAT_ENTITIES] == n_worlds * E and st[abi.STAT_WORLDS] == n_worlds
    assert st[abi.STAT_MASK_SUM] == after["vis_mask"].astype(np.uint64).sum()
    if dag == "chain3":
        return
    wa = np.moveaxis(f64(after["waabb"]), 1, 2)
    assert close(wa[..., :3], c).all() and close(wa[..., 3:], h, scale=1.0).all(), "world_aabb"
    check_bounds(after["wbounds"], bounds_of(c, h))
    g = after["gstats"].reshape(-1)
    assert g[abi.GSTAT_VISIBLE_LO:abi.GSTAT_VISIBLE_LO + 2].view(np.uint64)[0] == vis.sum()
    assert g[abi.GSTAT_WORLDS_LO:abi.GSTAT_WORLDS_LO + 2].view(np.uint64)[0] == n_worlds
    b = f64(g[abi.GSTAT_BOUNDS:abi.GSTAT_BOUNDS + 6].view(np.float32))
    per_world = bounds_of(c, h)
    assert close(b, np.concatenate([per_world[:, :3].min(0), per_world[:, 3:].max(0)])).all(), "group_stats bounds"
    if dag != "dag16":
        return
    assert np.array_equal(after["prev_world"].view(np.uint32), before["world"].view(np.uint32)), "prev_copy"
    vis2 = check_cull(after, ("vis_mask2", "vis_list2", "vis_count2"), c, h, before["frustum2"], E, tally)
    assert after["stats2"].reshape(-1).view(np.uint64)[abi.STAT_VISIBLE] == vis2.sum()
    check_bounds(after["vbounds"], bounds_of(c, h, bits(after["vis_mask"], E)))  # bounds of what `cull` found visible
    d2 = ((c - f64(before["campos"])[:, None, :, 0]) ** 2).sum(-1)  # lod = #{i : |c' - cam|^2 >= lod_dist2[i]}
    for i in range(3):
        limit = f64(before["lod_dist2"])[:, i, 0][:, None]
        decisions_match(after["lod"][:, 0, :] > i, (d2 - limit) / limit, "lod", tally)
    assert after["lod"].max() <= 3
    speed = np.abs(vel).sum(1)  # sleep bit = |vx| + |vy| + |vz| < eps, on the clamped velocity
    decisions_match(~bits(after["sleep"], E), speed - f64(before["sleep_eps"])[:, 0], "sleep_flags", tally)


@pytest.mark.parametrize("W,E,D", [(64, 1024, 4), (3, 8192, 8), (5, 77, 3)])
@pytest.mark.parametrize("dag", ["chain3", "dag8", "dag16"])
def test_every_generation_of_the_oracle_matches_the_float64_model(oracle, dag, W, E, D):
    K = 16
    data = synth.generate(W, E, D, dag=dag)
    names = tuple(data.keys())
    rig = synth.WorldStepRig(oracle, W, E, D, data={k: v.copy() for k, v in data.items()}, dag=dag)
    tally = [0, 0]  # decisions within 1e-3 of their threshold, decisions in all
    try:
        before = {k: v.copy() for k, v in data.items()}
        for _ in range(K):
            rig.step(1)
            after = rig.download(names)
            check_step(dag, before, after, E, tally)
            before = after
    finally:
        rig.close()
    assert tally[0] < 1e-3 * tally[1], f"{tally[0]} of {tally[1]} decisions sit within 1e-3 of their threshold"


def test_the_model_tells_a_wrong_oracle_apart():
    """The checker itself: a transposed rotation, a flipped plane normal or a missing |.| must not pass."""
    W, E, D = 4, 256, 3
    data = synth.generate(W, E, D)
    pos = f64(data["pos"]) + f64(data["vel"]) * f64(data["dt"])
    A, t = propagate(pos, f64(data["rot"]), f64(data["scale"]), data["parent"], data["level_start"])
    At, _ = propagate(pos, f64(data["rot"]) * np.array([-1, -1, -1, 1.0])[None, :, None], f64(data["scale"]), data["parent"],
                      data["level_start"])  # conjugate quaternion = transposed rotation
    assert not close(At, A, scale=1.0).all()
    c = np.einsum("weij,wej->wei", A, np.moveaxis(f64(data["aabb_c"]), 1, 2)) + t
    h = np.einsum("weij,wej->wei", np.abs(A), np.moveaxis(f64(data["aabb_h"]), 1, 2))
    good = slack(c, h, data["frustum"]) >= 0
    flipped = data["frustum"].copy()
    flipped[:, 0:3] *= -1  # first plane's normal
    assert (good != (slack(c, h, flipped) >= 0)).mean() > 0.01
    assert 0.02 < good.mean() < 0.98  # the synthetic frusta cut through the data: the plane test is exercised both ways
