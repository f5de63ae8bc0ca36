"""Synthetic worlds of the shape BASELINE.json names, and the rig that registers them on an
implementation of include/fimo_worlds_b200.h exactly as a user module would on the reference:
`resource_register` / `world_add_resource` / `system_register` / `system_group_add_systems` /
`system_group_schedule` (pkgs/fimo_worlds_meta/src/{resources,systems,worlds}.zig).

Harness-side only (tests, bench.py, smoke).  The generator is counter-based — every value is a pure
function of (seed, global world index, entity, field) — so any shard of any GPU materialises exactly
the data a single process would.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass

import numpy as np

from . import abi
from .worlds import Fence, Resource, System, World, create_pool

SEED = 0x46494D4F  # "FIMO"
DT = np.float32(1.0 / 60.0)

# name -> (scope, components, dtype); `level_start` gets depth+1 components at build time
LAYOUT = {
    "pos": (abi.SCOPE_ENTITY, 3, np.float32),
    "vel": (abi.SCOPE_ENTITY, 3, np.float32),
    "dt": (abi.SCOPE_WORLD, 1, np.float32),
    "rot": (abi.SCOPE_ENTITY, 4, np.float32),
    "scale": (abi.SCOPE_ENTITY, 1, np.float32),
    "parent": (abi.SCOPE_ENTITY, 1, np.uint32),
    "level_start": (abi.SCOPE_WORLD, 0, np.uint32),
    "world": (abi.SCOPE_ENTITY, 12, np.float32),
    "aabb_c": (abi.SCOPE_ENTITY, 3, np.float32),
    "aabb_h": (abi.SCOPE_ENTITY, 3, np.float32),
    "frustum": (abi.SCOPE_WORLD, 24, np.float32),
    "vis_mask": (abi.SCOPE_ENTITY_BITS, 1, np.uint32),
    "vis_list": (abi.SCOPE_ENTITY, 1, np.uint32),
    "vis_count": (abi.SCOPE_WORLD, 1, np.uint32),
    "stats": (abi.SCOPE_BATCH, 2 * abi.STAT_WORDS, np.uint32),
}
# resources the 8-system DAG adds (BASELINE.json configs[2]) ...
LAYOUT_DAG8 = {
    "gravity": (abi.SCOPE_WORLD, 3, np.float32),
    "angvel": (abi.SCOPE_ENTITY, 3, np.float32),
    "waabb": (abi.SCOPE_ENTITY, 6, np.float32),
    "wbounds": (abi.SCOPE_WORLD, 6, np.float32),
    "gstats": (abi.SCOPE_BATCH, abi.GSTAT_WORDS, np.uint32),
}
# ... and the 16-system DAG (configs[4])
LAYOUT_DAG16 = {
    "frustum2": (abi.SCOPE_WORLD, 24, np.float32),
    "vis_mask2": (abi.SCOPE_ENTITY_BITS, 1, np.uint32),
    "vis_list2": (abi.SCOPE_ENTITY, 1, np.uint32),
    "vis_count2": (abi.SCOPE_WORLD, 1, np.uint32),
    "stats2": (abi.SCOPE_BATCH, 2 * abi.STAT_WORDS, np.uint32),
    "prev_world": (abi.SCOPE_ENTITY, 12, np.float32),
    "lod": (abi.SCOPE_ENTITY, 1, np.uint32),
    "campos": (abi.SCOPE_WORLD, 3, np.float32),
    "lod_dist2": (abi.SCOPE_WORLD, 3, np.float32),
    "vmax": (abi.SCOPE_WORLD, 1, np.float32),
    "sleep": (abi.SCOPE_ENTITY_BITS, 1, np.uint32),
    "sleep_eps": (abi.SCOPE_WORLD, 1, np.float32),
    "damp": (abi.SCOPE_WORLD, 1, np.float32),
    "vbounds": (abi.SCOPE_WORLD, 6, np.float32),
    "gstats2": (abi.SCOPE_BATCH, abi.GSTAT_WORDS, np.uint32),
}
DAGS = ("chain3", "dag8", "dag16")


def layout_for(dag: str) -> dict:
    out = dict(LAYOUT)
    if dag in ("dag8", "dag16"):
        out.update(LAYOUT_DAG8)
    if dag == "dag16":
        out.update(LAYOUT_DAG16)
    return out


def outputs_for(dag: str) -> tuple:
    """Every resource some system of the DAG writes."""
    out = OUTPUTS
    if dag in ("dag8", "dag16"):
        out = out + ("vel", "rot", "waabb", "wbounds", "gstats")
    if dag == "dag16":
        out = out + ("angvel", "vis_mask2", "vis_list2", "vis_count2", "stats2", "prev_world", "lod", "sleep", "vbounds", "gstats2")
    return out
LIVE_IN = ("pos", "vel", "rot", "scale", "parent", "aabb_c", "aabb_h")  # 72 B / entity
LIVE_OUT = ("pos", "world", "vis_mask", "vis_list")  # 60.125 + 4 f_vis B / entity
OUTPUTS = ("pos", "world", "vis_mask", "vis_list", "vis_count", "stats")

_F = {  # field ids of the generator
    "pos": 0, "vel": 3, "rot": 6, "scale": 10, "aabb_c": 11, "aabb_h": 14, "parent": 17, "yaw": 18, "angvel": 19, "yaw2": 22,
    "campos": 23, "rot_retry": 32,
}


def pcg_hash(x: np.ndarray) -> np.ndarray:
    """PCG-RXS-M-XS 32 (one round), vectorised over uint32."""
    x = x.astype(np.uint32, copy=False)
    with np.errstate(over="ignore"):
        state = x * np.uint32(747796405) + np.uint32(2891336453)
        word = ((state >> ((state >> np.uint32(28)) + np.uint32(4))) ^ state) * np.uint32(277803737)
        return (word >> np.uint32(22)) ^ word


def _key(seed: int, world: np.ndarray, entity: np.ndarray, field: int) -> np.ndarray:
    with np.errstate(over="ignore"):
        return (np.uint32(seed) ^ (world.astype(np.uint32) * np.uint32(0x9E3779B9))
                ^ (entity.astype(np.uint32) * np.uint32(0x85EBCA6B)) ^ np.uint32((field * 0xC2B2AE35) & 0xFFFFFFFF))


def _uniform(seed, world, entity, field, lo, hi) -> np.ndarray:
    u = (pcg_hash(_key(seed, world, entity, field)) >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / (1 << 24))
    return (np.float32(lo) + np.float32(hi - lo) * u).astype(np.float32)


def level_starts(entities: int, depth: int) -> np.ndarray:
    """Level sizes proportional to 3^l (depth <= 4) or 2^l (deeper), remainder to the last level."""
    base = 3 if depth <= 4 else 2
    weights = [base ** l for l in range(depth)]
    total = sum(weights)
    sizes = [max(1, entities * w // total) for w in weights] if entities >= depth else [0] * depth
    if entities >= depth:
        sizes[-1] = entities - sum(sizes[:-1])
        assert sizes[-1] >= 1
    else:  # degenerate tiny worlds: everything is a root
        sizes[0] = entities
    starts = np.zeros(depth + 1, dtype=np.uint32)
    starts[1:] = np.cumsum(sizes)
    return starts


def generate(n_worlds: int, entities: int, depth: int, world0: int = 0, seed: int = SEED, dag: str = "chain3") -> dict[str, np.ndarray]:
    """All resources of `n_worlds` worlds starting at global world index `world0`, each laid out
    [n_worlds][components][count] (4-byte words), outputs zero-filled."""
    W, E = n_worlds, entities
    wi = (np.arange(W, dtype=np.uint32) + np.uint32(world0))[:, None]
    ei = np.arange(E, dtype=np.uint32)[None, :]
    out: dict[str, np.ndarray] = {}

    def field3(name, lo, hi, n=3):
        a = np.empty((W, n, E), dtype=np.float32)
        for k in range(n):
            a[:, k, :] = _uniform(seed, wi, ei, _F[name] + k, lo, hi)
        return a

    out["pos"] = field3("pos", -100.0, 100.0)
    out["vel"] = field3("vel", -1.0, 1.0)
    rot = field3("rot", -1.0, 1.0, 4)
    for attempt in range(1, 8):  # re-draw degenerate quaternions, deterministically
        bad = (rot.astype(np.float64) ** 2).sum(axis=1) < 1e-3
        if not bad.any():
            break
        wb, eb = np.nonzero(bad)
        for k in range(4):
            rot[wb, k, eb] = _uniform(seed, wi[wb, 0], ei[0, eb], _F["rot_retry"] + 4 * attempt + k, -1.0, 1.0)
    norm = np.sqrt((rot * rot).sum(axis=1, dtype=np.float32)).astype(np.float32)
    out["rot"] = (rot / norm[:, None, :]).astype(np.float32)
    out["scale"] = _uniform(seed, wi, ei, _F["scale"], 0.75, 1.25).reshape(W, 1, E)
    out["aabb_c"] = field3("aabb_c", -0.5, 0.5)
    out["aabb_h"] = field3("aabb_h", 0.1, 2.0)
    out["dt"] = np.full((W, 1, 1), DT, dtype=np.float32)

    starts = level_starts(E, depth)
    out["level_start"] = np.broadcast_to(starts[None, :, None], (W, depth + 1, 1)).copy()
    parent = np.full((W, E), abi.PARENT_ROOT, dtype=np.uint32)
    h = pcg_hash(_key(seed, wi, ei, _F["parent"]))
    for l in range(1, depth):
        lo, hi = int(starts[l]), int(starts[l + 1])
        n_prev = int(starts[l]) - int(starts[l - 1])
        if hi > lo and n_prev > 0:
            parent[:, lo:hi] = np.uint32(starts[l - 1]) + h[:, lo:hi] % np.uint32(n_prev)
    out["parent"] = parent.reshape(W, 1, E)

    # frustum: symmetric perspective, fov 90 deg, aspect 1, near 0.1, far 400, at the origin, yawed per world
    u = (pcg_hash(_key(seed, wi[:, 0], np.zeros(W, np.uint32), _F["yaw"])) >> np.uint32(8)).astype(np.float64) / (1 << 24)
    yaw = 2.0 * math.pi * u
    f = np.stack([np.sin(yaw), np.zeros(W), np.cos(yaw)], axis=1)
    r = np.stack([np.cos(yaw), np.zeros(W), -np.sin(yaw)], axis=1)
    up = np.broadcast_to(np.array([0.0, 1.0, 0.0]), (W, 3))
    planes = np.zeros((W, 6, 4), dtype=np.float64)

    def unit(v):
        return v / np.linalg.norm(v, axis=1, keepdims=True)

    def frustum_planes(f, r):
        planes[:, 0, :3], planes[:, 0, 3] = f, -0.1  # near
        planes[:, 1, :3], planes[:, 1, 3] = -f, 400.0  # far
        planes[:, 2, :3] = unit(f + r)  # left
        planes[:, 3, :3] = unit(f - r)  # right
        planes[:, 4, :3] = unit(f + up)  # bottom
        planes[:, 5, :3] = unit(f - up)  # top
        return planes.reshape(W, 24, 1).astype(np.float32)

    out["frustum"] = frustum_planes(f, r)

    out["world"] = np.zeros((W, 12, E), dtype=np.float32)
    out["vis_mask"] = np.zeros((W, 1, (E + 31) // 32), dtype=np.uint32)
    out["vis_list"] = np.zeros((W, 1, E), dtype=np.uint32)
    out["vis_count"] = np.zeros((W, 1, 1), dtype=np.uint32)
    out["stats"] = np.zeros((1, 2 * abi.STAT_WORDS, 1), dtype=np.uint32)

    if dag in ("dag8", "dag16"):
        out["gravity"] = np.broadcast_to(np.array([0.0, -9.81, 0.0], dtype=np.float32)[None, :, None], (W, 3, 1)).copy()
        out["angvel"] = field3("angvel", -2.0, 2.0)
        out["waabb"] = np.zeros((W, 6, E), dtype=np.float32)
        out["wbounds"] = np.zeros((W, 6, 1), dtype=np.float32)
        out["gstats"] = np.zeros((1, abi.GSTAT_WORDS, 1), dtype=np.uint32)
    if dag == "dag16":
        u2 = (pcg_hash(_key(seed, wi[:, 0], np.zeros(W, np.uint32), _F["yaw2"])) >> np.uint32(8)).astype(np.float64) / (1 << 24)
        yaw2 = 2.0 * math.pi * u2
        out["frustum2"] = frustum_planes(np.stack([np.sin(yaw2), np.zeros(W), np.cos(yaw2)], axis=1),
                                         np.stack([np.cos(yaw2), np.zeros(W), -np.sin(yaw2)], axis=1))
        out["vis_mask2"] = np.zeros((W, 1, (E + 31) // 32), dtype=np.uint32)
        out["vis_list2"] = np.zeros((W, 1, E), dtype=np.uint32)
        out["vis_count2"] = np.zeros((W, 1, 1), dtype=np.uint32)
        out["stats2"] = np.zeros((1, 2 * abi.STAT_WORDS, 1), dtype=np.uint32)
        out["prev_world"] = np.zeros((W, 12, E), dtype=np.float32)
        out["lod"] = np.zeros((W, 1, E), dtype=np.uint32)
        cam = np.empty((W, 3, 1), dtype=np.float32)
        for k in range(3):
            cam[:, k, 0] = _uniform(seed, wi[:, 0], np.zeros(W, np.uint32), _F["campos"] + 8 * k, -20.0, 20.0)
        out["campos"] = cam
        out["lod_dist2"] = np.broadcast_to(np.array([50.0 ** 2, 100.0 ** 2, 150.0 ** 2], dtype=np.float32)[None, :, None], (W, 3, 1)).copy()
        out["vmax"] = np.full((W, 1, 1), 0.9, dtype=np.float32)
        out["sleep"] = np.zeros((W, 1, (E + 31) // 32), dtype=np.uint32)
        out["sleep_eps"] = np.full((W, 1, 1), 0.75, dtype=np.float32)
        out["damp"] = np.full((W, 1, 1), 0.99, dtype=np.float32)
        out["vbounds"] = np.zeros((W, 6, 1), dtype=np.float32)
        out["gstats2"] = np.zeros((1, abi.GSTAT_WORDS, 1), dtype=np.uint32)
    return {k: np.ascontiguousarray(v) for k, v in out.items()}


def bytes_per_entity_step(f_vis: float) -> float:
    """B3 of SURVEY.md §8d: live-in read once (72 B) + live-out written once (60.125 + 4 f_vis B)."""
    return 72.0 + 60.125 + 4.0 * f_vis


def dag_bytes_per_entity_step(dag: str, f_vis: float, f_vis2: float = 0.0) -> float:
    """Algorithmic bytes per entity-step of a DAG by the same rule: every live-in resource read
    once, every live-out resource written once, per generation.
    dag8 : reads vel 12, pos 12, rot 16, angvel 12, scale 4, parent 4, aabb 24 = 84;
           writes vel 12, pos 12, rot 16, world 48, waabb 24, mask 0.125, list 4 f_vis.
    dag16: also reads world 48 (prev_copy, before propagate overwrites it); also writes angvel 12, prev_world 48,
           lod 4, sleep 0.125, mask2 0.125, list2 4 f_vis2."""
    if dag == "chain3":
        return bytes_per_entity_step(f_vis)
    b = 84.0 + 112.125 + 4.0 * f_vis
    if dag == "dag16":
        b += 48.0 + 12.0 + 48.0 + 4.0 + 0.125 + 0.125 + 4.0 * f_vis2
    return b


def integrate_matches(pos0: np.ndarray, vel: np.ndarray, steps: int, pos: np.ndarray) -> bool:
    """pos == `steps` x fma(vel, dt, pos0) in binary32.  The product of two binary32 values is exact in
    binary64, so one step is round32(round64(v*dt + p)); the double rounding can differ from a true fused
    multiply-add only in the last bit, which is what is allowed here (per step)."""
    lo = pos0.astype(np.float32).copy()
    hi = lo.copy()
    step = vel.astype(np.float64) * np.float64(DT)
    for _ in range(steps):
        a = (lo.astype(np.float64) + step).astype(np.float32)
        b = (hi.astype(np.float64) + step).astype(np.float32)
        lo = np.minimum(np.nextafter(a, np.float32(-np.inf)), np.nextafter(b, np.float32(-np.inf)))
        hi = np.maximum(np.nextafter(a, np.float32(np.inf)), np.nextafter(b, np.float32(np.inf)))
    return bool(((pos >= lo) & (pos <= hi)).all())


@dataclass
class RigOptions:
    fused: bool = True  # let the DAG compiler fuse integrate -> propagate -> cull into one world-step kernel
    graph: bool = True  # CUDA graph executor (False: plain stream launches with fork/join events)
    device: int = -1
    keep_host_copy: bool = True
    stats_ring: int = 0        # device pointer of a [stats_ring_slots][8] u64 ring for per-generation statistics snapshots
    stats_ring_slots: int = 0


class WorldStepRig:
    """One batch world + one system group running the 3-system chain
    integrate -> transform propagate -> AABB frustum cull on `lib`.

    On the product the resources are HBM SoA buffers behind FimoB200_ResourceView blobs and the
    systems are the built-in device systems; on the CPU oracle (tests / bench baseline only) the same
    descriptors are registered, the view's host_ptr is the data and the systems are ordinary
    `system_run` callbacks dispatched on the pool — the reference's CPU path."""

    def __init__(self, lib: abi.WorldsLib, n_worlds: int, entities: int, depth: int, world0: int = 0,
                 options: RigOptions | None = None, data: dict[str, np.ndarray] | None = None, pool: abi.Pool | None = None,
                 post_step=None, dag: str = "chain3", host_systems: bool = False, wind: tuple | None = None):
        """wind = (path of examples/wind_system/libwind_system.so, "between" | "front"): also registers that library's
        third-party system (X: vel, S: pos, wind, dt; a device entry point with builtin_id 0 on the product, a plain
        `system_run` on a CPU implementation) in the 8-system DAG — between `forces` and `integrate_lin`, or in front
        of `forces`."""
        assert dag in DAGS and not (host_systems and dag != "chain3") and not (wind and dag != "dag8")
        self.lib, self.W, self.E, self.depth, self.dag = lib, n_worlds, entities, depth, dag
        self.opt = options or RigOptions()
        self.on_device = lib.is_product
        self.host = data if data is not None else generate(n_worlds, entities, depth, world0, dag=dag)
        self._keep: list = []
        self._registered: list = []
        self.resources: dict[str, Resource] = {}
        self.views: dict[str, abi.ResourceView] = {}

        batch = None
        if self.on_device:
            batch = abi.WorldBatchExt(abi.ExtHeader(abi.STYPE_WORLD_BATCH, 0, None), n_worlds, entities, self.opt.device, 0)
        self.world = World.init(lib, "synthetic-batch", pool=pool, batch=batch)
        layout = layout_for(dag)
        if wind:
            layout["wind"] = (abi.SCOPE_WORLD, 4, np.float32)
            if "wind" not in self.host:  # direction xyz per world (counter-based like everything else), shear
                wi = (np.arange(n_worlds, dtype=np.uint32) + np.uint32(world0))
                w4 = np.empty((n_worlds, 4, 1), dtype=np.float32)
                for k in range(3):
                    w4[:, k, 0] = _uniform(SEED, wi, np.zeros(n_worlds, np.uint32), 40 + k, -3.0, 3.0)
                w4[:, 3, 0] = np.float32(0.01)
                self.host["wind"] = w4
        for name, (scope, comps, _dtype) in layout.items():
            if name == "level_start":
                comps = depth + 1
            arr = self.host[name]
            assert arr.dtype.itemsize == 4 and arr.flags["C_CONTIGUOUS"]
            ext = abi.ResourceDeviceExt(abi.ExtHeader(abi.STYPE_RESOURCE_DEVICE, 0, None), scope, comps, 4, 0)
            res = Resource.register(lib, name, C.sizeof(abi.ResourceView), 8,
                                    next_ptr=C.cast(C.pointer(ext), C.c_void_p), keep=(ext,))
            count = arr.shape[2]
            view = abi.ResourceView(None, arr.ctypes.data, arr.shape[0], count, comps, 4, 4 * count, 4 * comps * count, scope, 0)
            res.add_to_world(self.world, view)
            self.resources[name] = res
            self.views[name] = view

        gext = None
        if self.on_device:
            flags = (0 if self.opt.fused else abi.GROUP_FLAG_NO_FUSION) | (0 if self.opt.graph else abi.GROUP_FLAG_NO_GRAPH)
            self._post_step = abi.PostStepFn(post_step) if post_step is not None else abi.PostStepFn()
            if self.opt.stats_ring and self.opt.stats_ring_slots:
                flags |= abi.GROUP_FLAG_STATS_RING
            gext = abi.GroupDeviceExt(abi.ExtHeader(abi.STYPE_GROUP_DEVICE, 0, None), flags, 0, self._post_step, None,
                                      self.opt.stats_ring or None, self.opt.stats_ring_slots, 0)
        self.group = self.world.add_system_group("world-step", pool=pool, ext=gext)

        R = self.resources
        sysdef = self._system
        if dag == "chain3":
            integrate = sysdef("integrate", abi.BUILTIN_INTEGRATE, [R["pos"]], [R["vel"], R["dt"]])
            propagate = sysdef("propagate", abi.BUILTIN_PROPAGATE, [R["world"]],
                               [R["pos"], R["rot"], R["scale"], R["parent"], R["level_start"]], before=[integrate])
            cull = sysdef("cull", abi.BUILTIN_CULL, [R["vis_mask"], R["vis_list"], R["vis_count"], R["stats"]],
                          [R["world"], R["aabb_c"], R["aabb_h"], R["frustum"]], before=[propagate])
            self.systems = [integrate, propagate, cull]
            if host_systems:
                self.systems += self._host_systems(cull)
        else:
            # 8-system fan-out / fan-in DAG: forces -> integrate_lin || integrate_ang -> propagate ->
            # world_aabb -> { cull || bounds_reduce } -> group_stats
            wind_sys = self._wind_system(wind[0]) if wind and wind[1] == "front" else None
            forces = sysdef("forces", abi.BUILTIN_FORCES, [R["vel"]], [R["gravity"], R["dt"]], before=[wind_sys] if wind_sys else ())
            if wind and wind[1] == "between":
                wind_sys = self._wind_system(wind[0], before=[forces])
            integrate = sysdef("integrate_lin", abi.BUILTIN_INTEGRATE, [R["pos"]], [R["vel"], R["dt"]],
                               before=[forces, wind_sys] if wind and wind[1] == "between" else [forces])
            integrate_ang = sysdef("integrate_ang", abi.BUILTIN_INTEGRATE_ANG, [R["rot"]], [R["angvel"], R["dt"]])
            propagate = sysdef("propagate", abi.BUILTIN_PROPAGATE, [R["world"]],
                               [R["pos"], R["rot"], R["scale"], R["parent"], R["level_start"]], before=[integrate, integrate_ang])
            world_aabb = sysdef("world_aabb", abi.BUILTIN_WORLD_AABB, [R["waabb"]], [R["world"], R["aabb_c"], R["aabb_h"]],
                                before=[propagate])
            cull = sysdef("cull", abi.BUILTIN_CULL_WAABB, [R["vis_mask"], R["vis_list"], R["vis_count"], R["stats"]],
                          [R["waabb"], R["frustum"]], before=[world_aabb])
            bounds = sysdef("bounds_reduce", abi.BUILTIN_BOUNDS_REDUCE, [R["wbounds"]], [R["waabb"]], before=[world_aabb])
            gstats = sysdef("group_stats", abi.BUILTIN_GROUP_STATS, [R["gstats"]], [R["vis_count"], R["wbounds"]],
                            before=[cull, bounds])
            self.systems = [forces, integrate, integrate_ang, propagate, world_aabb, cull, bounds, gstats]
            if wind_sys is not None:
                self.systems.insert(0 if wind[1] == "front" else 1, wind_sys)
            if dag == "dag16":
                # 8 more systems; where no `before` is given the order follows from resource conflicts alone
                # (W/W, R/W, W/R in rank-then-add order, SystemGroup.zig:264-336):
                #   vel_clamp   W/W with forces on vel (after it: same rank, added later), read by integrate_lin
                #   angvel_damp writes angvel, read by integrate_ang
                #   prev_copy   reads world before propagate overwrites it (R/W)
                #   sleep_flags reads the clamped vel
                vel_clamp = sysdef("vel_clamp", abi.BUILTIN_VEL_CLAMP, [R["vel"]], [R["vmax"]])
                angvel_damp = sysdef("angvel_damp", abi.BUILTIN_ANGVEL_DAMP, [R["angvel"]], [R["damp"]])
                prev_copy = sysdef("prev_copy", abi.BUILTIN_PREV_COPY, [R["prev_world"]], [R["world"]])
                sleep = sysdef("sleep_flags", abi.BUILTIN_SLEEP_FLAGS, [R["sleep"]], [R["vel"], R["sleep_eps"]])
                cull2 = sysdef("cull2", abi.BUILTIN_CULL_WAABB, [R["vis_mask2"], R["vis_list2"], R["vis_count2"], R["stats2"]],
                               [R["waabb"], R["frustum2"]], before=[world_aabb])
                lod = sysdef("lod", abi.BUILTIN_LOD, [R["lod"]], [R["waabb"], R["campos"], R["lod_dist2"]], before=[world_aabb])
                vis_bounds = sysdef("vis_bounds", abi.BUILTIN_VIS_BOUNDS, [R["vbounds"]], [R["waabb"], R["vis_mask"]], before=[cull])
                gstats2 = sysdef("group_stats2", abi.BUILTIN_GROUP_STATS, [R["gstats2"]], [R["vis_count2"], R["vbounds"]],
                                 before=[cull2, vis_bounds])
                self.systems = [forces, vel_clamp, integrate, angvel_damp, integrate_ang, prev_copy, propagate, world_aabb, cull,
                                cull2, bounds, vis_bounds, lod, sleep, gstats, gstats2]
        self.group.add_systems(self.systems)
        self._fence = Fence(lib)
        self.closed = False

    def _host_systems(self, cull):
        """Two ordinary host systems (`system_run` callbacks on the pool) in the same group as the device systems
        (SURVEY.md §8 f2): `host_camera` writes the frusta every generation (X: frustum — ordered before `cull` by the
        W/R conflict alone), `host_census` reads the per-world visible counts after `cull` (S: vis_count, explicit
        dependency) into a plain host resource.  Both see device resources as `ResourceView.host_ptr`: the data
        itself on the CPU oracle, a coherent pinned mirror on the product."""
        class Census(C.Structure):
            _fields_ = [("visible", C.c_uint64), ("generations", C.c_uint64), ("last", C.c_uint64)]

        self.Census = Census
        census = Resource.register(self.lib, "census", C.sizeof(Census), 8)
        census.add_to_world(self.world, Census(0, 0, 0))
        self.resources["census"] = census
        self.host_errors: list = []

        def words(blob_ptr, dtype):
            view = C.cast(blob_ptr, C.POINTER(abi.ResourceView)).contents
            n = view.n_worlds * view.components * view.count
            return np.ctypeslib.as_array(C.cast(view.host_ptr, C.POINTER(C.c_uint32)), shape=(n,)).view(dtype).reshape(
                view.n_worlds, view.components, view.count)

        def run_camera(_ctx, ex, _sh, _fence):
            try:
                fr = words(ex[0], np.float32)
                fr[...] = np.roll(fr, 1, axis=0)  # every world takes its neighbour's camera
            except Exception as e:  # noqa: BLE001 - reported by the test, never swallowed
                self.host_errors.append(e)

        def run_census(_ctx, ex, sh, _fence):
            try:
                c = C.cast(ex[0], C.POINTER(Census)).contents
                last = int(words(sh[0], np.uint32).sum())
                c.visible += last
                c.generations += 1
                c.last = last
            except Exception as e:  # noqa: BLE001
                self.host_errors.append(e)

        R = self.resources
        camera = System.register(self.lib, "host_camera", run=run_camera, exclusive=[R["frustum"]])
        counted = System.register(self.lib, "host_census", run=run_census, exclusive=[census], shared=[R["vis_count"]], before=[cull])
        self._registered += [camera, counted]
        return [camera, counted]

    def _wind_system(self, path: str, before=()):
        """The third-party system of examples/wind_system: its library fills the descriptor (factory, state size,
        create / destroy / run) and, on the product, the device extension with ITS entry point (builtin_id 0)."""
        windlib = C.CDLL(path)
        windlib.wind_fill_descriptor.argtypes = [C.c_void_p, C.c_void_p, C.c_float]
        windlib.wind_fill_descriptor.restype = None
        windlib.wind_counters.argtypes = [C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        windlib.wind_bind_futex_wake.argtypes = [C.c_void_p]
        windlib.wind_bind_futex_wake(C.cast(self.lib.fimo_tasks_futex_wake, C.c_void_p))
        desc = abi.SystemDescriptor()
        ext = abi.SystemDeviceExt() if self.on_device else None
        windlib.wind_fill_descriptor(C.byref(desc), C.byref(ext) if ext is not None else None, 1.25)
        self.windlib = windlib
        R = self.resources
        system = System.register(self.lib, "wind", exclusive=[R["vel"]], shared=[R["pos"], R["wind"], R["dt"]], before=before,
                                 descriptor=desc, keep=[desc, ext, windlib])
        self._registered.append(system)
        return system

    def wind_counters(self) -> tuple[int, int]:
        d, h = C.c_uint64(0), C.c_uint64(0)
        self.windlib.wind_counters(C.byref(d), C.byref(h))
        return int(d.value), int(h.value)

    def census(self):
        """The host resource `host_census` writes (locked shared, copied out)."""
        ptr = self.resources["census"].lock_in_world(self.world, False)
        try:
            c = C.cast(ptr, C.POINTER(self.Census)).contents
            return int(c.visible), int(c.generations), int(c.last)
        finally:
            self.resources["census"].unlock_in_world(self.world, False)

    def _system(self, label, builtin, exclusive, shared, before=()):
        desc = abi.SystemDescriptor()
        keep = [desc]
        if self.on_device:
            ext = abi.SystemDeviceExt()
            self.lib.check(self.lib.fimo_b200_builtin_system(builtin, C.byref(desc), C.byref(ext)))
            keep.append(ext)
        else:
            self.lib.check(self.lib.fimo_oracle_builtin_system(builtin, C.byref(desc)))
        system = System.register(self.lib, label, exclusive=exclusive, shared=shared, before=before, descriptor=desc, keep=keep)
        self._registered.append(system)
        return system

    # ---- stepping -----------------------------------------------------------------------------------
    def schedule(self, steps: int = 1, signal_last: bool = True) -> None:
        """`steps` x system_group_schedule; the last one signals the rig's fence."""
        self._fence.reset()
        for i in range(steps):
            self.group.schedule(signal=self._fence if (signal_last and i == steps - 1) else None)

    def wait(self) -> None:
        self._fence.wait()

    def step(self, steps: int = 1) -> None:
        self.schedule(steps)
        self.wait()

    def executor_info(self) -> abi.ExecutorInfo:
        info = abi.ExecutorInfo()
        self.lib.check(self.lib.fimo_b200_group_executor_info(self.group.handle, C.byref(info)))
        return info

    def executor_text(self) -> str:
        """fimo_b200_group_executor_describe: kernels and launch geometry of the compiled executor."""
        self.executor_info()
        buf = C.create_string_buffer(512)
        self.lib.fimo_b200_group_executor_describe(self.group.handle, buf, 512)
        return buf.value.decode()

    def kernels_launched(self) -> int:
        return int(self.lib.fimo_b200_group_kernels_launched(self.group.handle))

    def device_ms(self) -> tuple[float, int]:
        n = C.c_uint64(0)
        ms = self.lib.fimo_b200_group_device_ms(self.group.handle, C.byref(n))
        return ms, n.value

    def span_ms(self) -> float:
        return self.lib.fimo_b200_group_span_ms(self.group.handle)

    # ---- host access through the ABI ----------------------------------------------------------------
    def _locked_view(self, name: str, exclusive: bool) -> abi.ResourceView:
        ptr = self.resources[name].lock_in_world(self.world, exclusive)
        return C.cast(ptr, C.POINTER(abi.ResourceView)).contents

    def download(self, names=None) -> dict[str, np.ndarray]:
        """world_lock_resources (shared) -> copy out -> unlock, per resource."""
        names = outputs_for(self.dag) if names is None else names
        out = {}
        for name in names:
            view = self._locked_view(name, exclusive=False)
            try:
                if self.on_device:
                    arr = np.empty_like(self.host[name])
                    self.lib.check(self.lib.fimo_b200_resource_download(C.byref(view), arr.ctypes.data))
                else:
                    arr = self.host[name].copy()
            finally:
                self.resources[name].unlock_in_world(self.world, False)
            out[name] = arr
        return out

    def download_into(self, name: str, arr: np.ndarray) -> None:
        """download() into a caller-provided (e.g. pinned) host buffer."""
        view = self._locked_view(name, exclusive=False)
        try:
            if self.on_device:
                self.lib.check(self.lib.fimo_b200_resource_download(C.byref(view), arr.ctypes.data))
            else:
                arr[...] = self.host[name]
        finally:
            self.resources[name].unlock_in_world(self.world, False)

    def upload(self, name: str, arr: np.ndarray) -> None:
        """world_lock_resources (exclusive) -> copy in -> unlock."""
        assert arr.dtype.itemsize == 4 and arr.flags["C_CONTIGUOUS"] and arr.shape == self.host[name].shape
        view = self._locked_view(name, exclusive=True)
        try:
            if self.on_device:
                self.lib.check(self.lib.fimo_b200_resource_upload(C.byref(view), arr.ctypes.data))
            else:
                self.host[name][...] = arr
        finally:
            self.resources[name].unlock_in_world(self.world, True)

    # ---- stream-ordered transfers (product only): queued like generations, no host lock dance ---------
    def upload_async(self, name: str, arr: np.ndarray) -> None:
        """H2D copy of a whole resource, ordered before every generation scheduled afterwards; `arr` (pinned
        for a truly asynchronous copy) must stay untouched until a later fence has been waited on."""
        assert arr.dtype.itemsize == 4 and arr.flags["C_CONTIGUOUS"] and arr.shape == self.host[name].shape
        self.lib.check(self.lib.fimo_b200_group_transfer(self.group.handle, self.resources[name].handle, arr.ctypes.data,
                                                         abi.TRANSFER_HOST_TO_DEVICE, None))

    def download_async(self, name: str, arr: np.ndarray, signal: Fence | None = None) -> None:
        """D2H copy of a whole resource, ordered after every generation scheduled before; `signal` fires when
        the data has landed in `arr`."""
        assert arr.dtype.itemsize == 4 and arr.flags["C_CONTIGUOUS"] and arr.shape == self.host[name].shape
        self.lib.check(self.lib.fimo_b200_group_transfer(self.group.handle, self.resources[name].handle, arr.ctypes.data,
                                                         abi.TRANSFER_DEVICE_TO_HOST, signal.ptr() if signal else None))

    def io_list(self, arrays: dict[str, np.ndarray]):
        """A FimoB200_StepTransfer array over whole-resource host buffers (build once, reuse every step)."""
        items = (abi.StepTransfer * max(1, len(arrays)))()
        for i, (name, arr) in enumerate(arrays.items()):
            assert arr.dtype.itemsize == 4 and arr.flags["C_CONTIGUOUS"] and arr.shape == self.host[name].shape
            items[i].resource = self.resources[name].handle
            items[i].host = arr.ctypes.data
        return items, len(arrays)

    def schedule_io(self, uploads, downloads, signal: Fence | None = None) -> None:
        """fimo_b200_group_schedule_io: one generation with its input copies in front and its result copies behind,
        one call, one completion; `uploads` / `downloads` come from io_list()."""
        self.lib.check(self.lib.fimo_b200_group_schedule_io(self.group.handle, uploads[0], uploads[1], downloads[0], downloads[1],
                                                            None, 0, signal.ptr() if signal else None))

    def readback_words(self, capacity_entries: int) -> int:
        """u32 words of a fimo_b200_group_schedule_readback host buffer whose list holds `capacity_entries`."""
        return abi.READBACK_HEADER_WORDS + 2 * self.W + self.W * ((self.E + 31) // 32) + capacity_entries

    def readback_desc(self, host: np.ndarray) -> abi.CullReadback:
        R = self.resources
        return abi.CullReadback(R["vis_mask"].handle, R["vis_list"].handle, R["vis_count"].handle, host.ctypes.data, host.nbytes)

    def schedule_readback(self, uploads, desc: abi.CullReadback, signal: Fence | None = None) -> None:
        """fimo_b200_group_schedule_readback: one generation whose packed cull outputs land in the host buffer of `desc`."""
        self.lib.check(self.lib.fimo_b200_group_schedule_readback(self.group.handle, uploads[0], uploads[1], C.byref(desc), None, 0,
                                                                  signal.ptr() if signal else None))

    def parse_readback(self, host: np.ndarray) -> dict:
        """Splits a landed read-back buffer into its parts (views of `host`)."""
        W, words = self.W, (self.E + 31) // 32
        h = host[:8]
        at = 8
        counts, offsets = host[at:at + W], host[at + W:at + 2 * W]
        at += 2 * W
        mask = host[at:at + W * words].reshape(W, words)
        at += W * words
        capacity = int(h[3])
        return {"total": int(h[0]), "n_worlds": int(h[1]), "mask_words": int(h[2]), "capacity": capacity, "truncated": int(h[4]),
                "generation": int(h[5]) | (int(h[6]) << 32), "entities": int(h[7]), "counts": counts, "offsets": offsets, "mask": mask,
                "list": host[at:at + min(int(h[0]), capacity)]}

    def stats(self) -> np.ndarray:
        return self.download(("stats",))["stats"].reshape(-1).view(np.uint64)

    def close(self) -> None:
        if self.closed:
            return
        self.closed = True
        for s in self.systems:
            self.group.remove_system(s)
        for s in reversed(self._registered):  # dependants first
            s.unregister()
        self.group.deinit()
        for name, res in self.resources.items():
            out = abi.ResourceView()
            res.remove_from_world(self.world, out)
            res.unregister()
        self.world.deinit()


BATCH_SCOPE = ("stats", "stats2", "gstats", "gstats2")  # one instance per batch: merged when a batch is stepped in slices


def _merge_batch_words(name: str, parts: list[np.ndarray]) -> np.ndarray:
    """Batch-scope words of a batch stepped as several sub-batches: every FIMO_B200_STAT_* word is an additive
    count; FIMO_B200_GSTAT_*: two additive u64 counts and min / max bounds (exact in any order)."""
    if name.startswith("stats"):
        return np.sum([p.reshape(-1).view(np.uint64) for p in parts], axis=0, dtype=np.uint64).view(np.uint32).reshape(parts[0].shape)
    out = np.zeros(abi.GSTAT_WORDS, dtype=np.uint32)
    flat = [p.reshape(-1) for p in parts]
    for lo in (abi.GSTAT_VISIBLE_LO, abi.GSTAT_WORLDS_LO):
        out[lo:lo + 2] = np.sum([f[lo:lo + 2].view(np.uint64) for f in flat], axis=0, dtype=np.uint64).view(np.uint32)
    b = abi.GSTAT_BOUNDS
    out[b:b + 3] = np.min([f[b:b + 3].view(np.float32) for f in flat], axis=0).view(np.uint32)
    out[b + 3:b + 6] = np.max([f[b + 3:b + 6].view(np.float32) for f in flat], axis=0).view(np.uint32)
    return out.reshape(parts[0].shape)


def step_in_slices(lib: abi.WorldsLib, data: dict[str, np.ndarray], entities: int, depth: int, steps: int, dag: str = "chain3",
                   slices: int = 0, workers: int = 0) -> dict[str, np.ndarray]:
    """Steps a whole batch on a CPU implementation (the oracle: tests / bench parity gate only) as `slices`
    contiguous sub-batches of worlds — one World + SystemGroup each, all on ONE shared pool, every generation of every
    slice queued before the first wait — so that a BASELINE-size batch takes seconds instead of minutes.  Worlds are
    independent (World.zig:26), so the result equals the one-batch run; `data` is stepped IN PLACE (each slice's
    resources are views of it) and the outputs are returned with the batch-scope words merged."""
    import os

    W = data["pos"].shape[0]
    workers = workers or (os.cpu_count() or 1)
    slices = max(1, min(slices or workers, W))
    pool = create_pool(lib, "sliced-batch", workers)
    bounds = [(i * W // slices, (i + 1) * W // slices) for i in range(slices)]
    rigs, locals_ = [], []
    try:
        for lo, hi in bounds:
            part = {k: (v[lo:hi] if k not in BATCH_SCOPE else v.copy()) for k, v in data.items()}
            locals_.append(part)
            rigs.append(WorldStepRig(lib, hi - lo, entities, depth, data=part, pool=pool, dag=dag))
        for r in rigs:
            r.schedule(steps)
        for r in rigs:
            r.wait()
        out = {}
        for name in outputs_for(dag):
            out[name] = _merge_batch_words(name, [p[name] for p in locals_]) if name in BATCH_SCOPE else data[name]
        return out
    finally:
        for r in rigs:
            r.close()
        pool.unref()


def compare_outputs(got: dict[str, np.ndarray], want: dict[str, np.ndarray], float_rtol: float = 0.0) -> list[str]:
    """Integer / index / mask outputs must match bit for bit; float outputs within `float_rtol`
    relative (0.0 = bit-exact, the design target; north_star allows 1e-5)."""
    problems = []
    for name, w in want.items():
        g = got[name]
        if g.shape != w.shape:
            problems.append(f"{name}: shape {g.shape} != {w.shape}")
            continue
        if w.dtype == np.float32 and float_rtol > 0.0:
            bad = ~np.isclose(g, w, rtol=float_rtol, atol=0.0, equal_nan=True)
        else:
            bad = g.view(np.uint32) != w.view(np.uint32)
        if bad.any():
            idx = tuple(int(i[0]) for i in np.nonzero(bad))
            problems.append(f"{name}: {int(bad.sum())} of {bad.size} words differ, first at {idx}: got {g[idx]!r} want {w[idx]!r}")
    return problems
