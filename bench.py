#!/usr/bin/env python
"""bench.py — batched world-steps/sec of the fimo_worlds system-group step on B200 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W]            the CUDA product, one process per GPU
  python bench.py --impl reference [...]                         the reference's CPU path (oracle restatement) on host cores

A "step" is one generation (`system_group_schedule`) of the 3-system chain integrate -> transform propagate
-> AABB frustum cull over one batch of synthetic worlds.  N=1 workload = BASELINE.json configs[1]:
4096 worlds x 1024 entities, depth-4 hierarchy.  N>1: the same batch per GPU (weak scaling), worlds sharded by
global index, no data-path collective; the end-of-step statistics are exchanged by the product itself
(fimo_b200_group_stats_peers_bind: peer stores over NVLink from the publishing CTA; NCCL all-reduce as the checked
second transport).  Prints ONE JSON line on rank 0 — and refuses to print it if the full-size parity gate (K = 2
generations of this very workload against the CPU oracle, bit for bit) fails.

Timing: W warm-up steps, then `--windows` windows of EXACTLY K steps each, every window bracketed by a barrier +
synchronize on both sides and timed on the device (CUDA events on the group's stream, first generation's start ->
last generation's end, max over ranks); `value` is taken from the MEDIAN window and min / max are reported beside it.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "batched world-steps/sec"
UNIT = "world-steps/s"
WORLDS_PER_GPU, ENTITIES, DEPTH = 4096, 1024, 4
CPU_SAMPLE_WORLDS = 256  # bounded sample of the same workload for the CPU arm
CPU_QUEUE_DEPTH = 32     # generations queued per group before the CPU arm waits (schedule is asynchronous)


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=200)
    p.add_argument("--warmup", type=int, default=20)
    p.add_argument("--windows", type=int, default=5, help="timed windows of --steps generations each; the median is reported")
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--worlds", type=int, default=WORLDS_PER_GPU, help="worlds per GPU")
    p.add_argument("--entities", type=int, default=ENTITIES)
    p.add_argument("--depth", type=int, default=DEPTH)
    p.add_argument("--dag", default="chain3", choices=["chain3", "dag8", "dag16"],
                   help="system set: the 3-system chain (configs[1]), the 8-system fan-out/fan-in DAG (configs[2]) or the "
                        "16-system DAG with mixed read/write conflicts (configs[4])")
    p.add_argument("--no-fusion", action="store_true", help="one kernel per system (CUDA graph of the DAG)")
    p.add_argument("--no-graph", action="store_true", help="plain stream launches instead of a CUDA graph per generation")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--no-parity-gate", action="store_true", help="skip the K=2 full-size check against the oracle (sweeps only)")
    p.add_argument("--no-other-configs", action="store_true", help="skip the short configs[2] / configs[3]-shape / configs[4]-shape measurements")
    p.add_argument("--stats", default="peers", choices=["peers", "nccl"], help="transport of the statistics exchange (N > 1)")
    p.add_argument("--stats-bucket", type=int, default=8, help="NCCL transport: generations per all-reduce")
    p.add_argument("--cpu-steps", type=int, default=0, help="steps of the CPU arm (0 = sized for ~15 s)")
    return p.parse_args()


def measured_peaks() -> tuple[float, str]:
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(dag: str, W: int, E: int, D: int):
    """DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture of THIS config, or None."""
    try:
        table = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        entry = table.get("configs", {}).get(f"{dag}:{W}x{E}:d{D}")
        return (entry["dram_bytes_per_launch"], entry["source"]) if entry else (None, None)
    except Exception:
        return None, None


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference's own structure — one World + one SystemGroup per world, every group scheduled
# on ONE shared pool with all host threads, systems as system_run callbacks.
# ------------------------------------------------------------------------------------------------
DAG_TEXT = {"chain3": "3-system chain (integrate -> propagate -> AABB frustum cull)",
            "dag8": "8-system fan-out/fan-in DAG (forces -> integrate_lin || integrate_ang -> propagate -> world_aabb -> cull || bounds -> stats)",
            "dag16": "16-system DAG with mixed read/write resource conflicts"}


def cpu_arm(worlds: int, entities: int, depth: int, steps: int, warmup: int, dag: str = "chain3", seconds_budget: float = 15.0) -> dict:
    from fimo_b200 import synth
    from fimo_b200.worlds import create_pool
    from oracle import loader

    lib = loader.load()
    cores = int(lib.fimo_oracle_hardware_threads()) or (os.cpu_count() or 1)
    pool = create_pool(lib, "cpu-baseline", cores)
    data = synth.generate(worlds, entities, depth, dag=dag)
    rigs = []
    for w in range(worlds):
        one = {k: (v if v.shape[0] != worlds else v[w:w + 1]).copy() for k, v in data.items()}
        rigs.append(synth.WorldStepRig(lib, 1, entities, depth, data=one, pool=pool, dag=dag))
    groups = (C.c_void_p * worlds)(*[r.group.handle for r in rigs])

    def run(n):  # generations queue behind each group's timeline semaphore, CPU_QUEUE_DEPTH at a time (SystemGroup.zig:430-452)
        t, left = 0.0, n
        while left > 0:
            k = min(left, CPU_QUEUE_DEPTH)
            t += lib.fimo_oracle_step_groups_queued(groups, worlds, k)
            left -= k
        return t

    if steps <= 0:  # size the run for ~seconds_budget of CPU work
        t = run(4)
        steps = max(4, min(4000, int(seconds_budget / max(t / 4, 1e-6))))
    run(max(warmup, 1))
    seconds = run(steps)
    for r in rigs:
        r.close()
    pool.unref()
    return {
        "value": worlds * steps / seconds, "unit": UNIT, "cores": cores, "kind": "port",
        "sample": f"{worlds} worlds x {entities} entities, depth {depth}, {dag}, {steps} steps, one World+SystemGroup "
                  f"per world on one shared {cores}-thread pool, {CPU_QUEUE_DEPTH} generations queued per group at a time (oracle "
                  f"restatement of fimo_worlds/fimo_tasks: one event-loop thread per pool, every task a stackful fiber on a cached 8 MiB stack; gcc -O3 -mavx2 -mfma; the reference itself needs zig + boost.context and cannot be built here)",
        "seconds": seconds, "steps": steps, "ms_per_step": 1e3 * seconds / steps,
    }


def config_of(worlds: int, entities: int, depth: int, dag: str) -> dict:
    """The `config` object of BOTH arms (the driver compares them key for key): what is stepped, nothing about how."""
    from fimo_b200 import synth

    floor_mb = synth.dag_bytes_per_entity_step(dag, 0.0, 0.0) * worlds * entities / 1e6  # without the visible lists
    return {"workload": f"{worlds} worlds x {entities} entities per GPU, depth-{depth} transform hierarchy, {DAG_TEXT[dag]}"
                        + (", BASELINE.json configs[1]" if (worlds, entities, depth, dag) == (4096, 1024, 4, "chain3") else ""),
            "dag": dag, "worlds_per_gpu": worlds, "entities_per_world": entities, "depth": depth,
            "l2": ("inputs larger than L2 (state touched per step >= %.0f MB > 126 MB), no flush between steps" % floor_mb) if floor_mb > 126.0 else
                  ("state touched per step >= %.0f MB (partly L2-resident between steps; no flush)" % floor_mb)}


def run_reference(args) -> None:
    """The reference arm: the SAME config, steps and warm-up as the B200 arm wherever that ends within a few minutes —
    every world of the batch gets its own World + SystemGroup, as on the reference — else (over 8 M entities) a
    bounded sample of the worlds, said so in the line."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    worlds = args.worlds if args.worlds * args.entities <= (8 << 20) else max(1, (8 << 20) // max(args.entities, 1))
    steps, warmup = args.steps, max(args.warmup, 3)
    base = cpu_arm(worlds, args.entities, args.depth, steps, warmup, args.dag)
    sampled = "" if worlds == args.worlds else f"; CPU arm steps a {worlds}-world sample per step"
    line = {
        "impl": "reference", "metric": METRIC, "value": base["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": base["steps"], "warmup": warmup, "ms_per_step": base["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_of(args.worlds, args.entities, args.depth, args.dag),
        "same_config": worlds == args.worlds, "sampled": sampled.lstrip("; ") or None,
        "cpu_baseline": {k: base[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# clocks sampled DURING the timed region
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock and throttle reasons of one GPU, sampled DURING the timed region: NVML in-process (a thread every 2 ms plus one
    explicit sample per timed window while its generations are in flight — a 2 ms window would otherwise fall between two
    `nvidia-smi -lms 100` lines), nvidia-smi as the fallback."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device: int):
        self.rows, self.proc, self.nvml, self.handle, self.stop = [], None, None, None, False
        try:
            import pynvml
            pynvml.nvmlInit()
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            index = int(visible.split(",")[device]) if visible and all(v.strip().isdigit() for v in visible.split(",")) else device
            try:  # NVML enumerates in PCI order, CUDA need not: match the device by its PCI address
                import torch
                pr = torch.cuda.get_device_properties(device)
                self.handle = pynvml.nvmlDeviceGetHandleByPciBusId(("%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)).encode())
            except Exception:
                self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.nvml = pynvml
            self.sm_max = int(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.sample()
            threading.Thread(target=self._poll, daemon=True).start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(device)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def sample(self) -> None:
        """One NVML sample now (called with a window's generations in flight)."""
        if self.nvml is None:
            return
        n = self.nvml
        try:
            sm = int(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM))
            bits = int(n.nvmlDeviceGetCurrentClocksEventReasons(self.handle)) if hasattr(n, "nvmlDeviceGetCurrentClocksEventReasons") \
                else int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
            flag = lambda name, old: bits & int(getattr(n, name, getattr(n, old, 0)))  # noqa: E731
            row = ["0", str(sm), str(self.sm_max), "", "",
                   "Active" if flag("nvmlClocksEventReasonHwSlowdown", "nvmlClocksThrottleReasonHwSlowdown") else "Not Active",
                   "Active" if flag("nvmlClocksEventReasonHwThermalSlowdown", "nvmlClocksThrottleReasonHwThermalSlowdown") else "Not Active",
                   "Active" if flag("nvmlClocksEventReasonSwThermalSlowdown", "nvmlClocksThrottleReasonSwThermalSlowdown") else "Not Active",
                   "Active" if flag("nvmlClocksEventReasonSwPowerCap", "nvmlClocksThrottleReasonSwPowerCap") else "Not Active"]
            self.rows.append((time.time(), row))
        except Exception:
            pass

    def _poll(self):
        while not self.stop:
            self.sample()
            time.sleep(0.002)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def summary(self, t0: float, t1: float) -> dict:
        self.stop = True
        if self.proc is not None:
            self.proc.terminate()
        rows = [r for t, r in self.rows if t0 <= t <= t1 + 0.15] or [r for _, r in self.rows[-3:]]
        sm = sorted(int(float(r[1])) for r in rows if len(r) > 2 and r[1].replace(".", "").isdigit())
        mx = [int(float(r[2])) for r in rows if len(r) > 2 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in rows if len(r) >= 9 for n, v in zip(names, r[5:9]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(rows), "source": "nvml" if self.nvml is not None else "nvidia-smi"}


# ------------------------------------------------------------------------------------------------
# the parity gate: the benched workload on the launch shape the timed windows use, product vs oracle, bit for bit
# ------------------------------------------------------------------------------------------------
def parity_gate(lib, data, W, E, D, dag, options, steps: int = 0) -> dict:
    """Gates the path that is timed.  The timed windows schedule K generations back to back, and when the plan lets queued
    generations share a launch (max_generations_per_launch > 1: grid over all worlds, block-of-worlds order, bulk-group
    waits across generation boundaries, one statistics publish for the whole launch) almost all of them run that way —
    two generations never do.  So the gate queues max_generations_per_launch + 4 generations without waiting, checks
    that launches really were shared (fewer launches than one per generation; retried on a fresh rig if the host was too
    slow to queue behind the first launches) and compares every written resource with the oracle."""
    from fimo_b200 import synth
    from oracle import loader

    t0 = time.time()
    shared, launches, tries = False, 0, 0
    while True:
        tries += 1
        rig = synth.WorldStepRig(lib, W, E, D, data=data, options=options, dag=dag)  # uploads `data`, never writes it
        try:
            info = rig.executor_info()
            per_launch, per_generation = int(info.max_generations_per_launch), int(info.kernel_launches_per_generation)
            K = steps or (per_launch + 4 if per_launch > 1 else 2)
            before = rig.kernels_launched()
            rig.step(K)  # K x schedule, ONE wait
            launches = rig.kernels_launched() - before
            got = rig.download()
        finally:
            rig.close()
        shared = launches < K * per_generation
        if shared or per_launch <= 1 or tries == 3:
            break
    want = synth.step_in_slices(loader.load(), {k: v.copy() for k, v in data.items()}, E, D, K, dag=dag)
    problems = synth.compare_outputs(got, want, 0.0)
    return {"ok": not problems, "problems": problems[:4], "steps": K, "seconds": round(time.time() - t0, 2),
            "kernel_launches": int(launches), "max_generations_per_launch": per_launch,
            "covers_shared_launches": bool(shared) if per_launch > 1 else None,
            "checked": f"{W} x {E}, {dag}: every written resource after {K} generations queued without waiting "
                       f"({launches} kernel launches), bit-exact vs the CPU oracle"}


# ------------------------------------------------------------------------------------------------
# device-resident measurement of one configuration
# ------------------------------------------------------------------------------------------------
def timed_windows(rig, K: int, windows: int, barrier, reduce_max=None, sampler=None):
    """`windows` x exactly K generations, each bracketed by barrier + synchronize; per window (span_ms, kernel_ms)."""
    spans, kernels, walls = [], [], []
    timed_windows.launches = []  # kernel launches inside each window, counted by the executor
    for _ in range(windows):
        rig.device_ms()
        rig.span_ms()
        barrier()
        launched0 = rig.kernels_launched()
        t0 = time.perf_counter()
        rig.schedule(K)
        if sampler is not None:
            sampler.sample()  # the window's generations are in flight
        rig.wait()
        barrier()
        walls.append(1e3 * (time.perf_counter() - t0))
        timed_windows.launches.append(rig.kernels_launched() - launched0)
        kernel_ms, timed = rig.device_ms()
        span_ms = rig.span_ms()
        spans.append(span_ms if (timed == K and span_ms > 0.0) else walls[-1])
        kernels.append(kernel_ms)
    if reduce_max is not None:
        spans = reduce_max(spans)
    return spans, kernels, walls


def roofline_of(rig, np, abi, synth, dag, W, E, D, kernel_ms, step_ms, peak, peak_src, info) -> dict:
    stats = rig.stats()
    f_vis = float(stats[abi.STAT_VISIBLE]) / max(float(stats[abi.STAT_ENTITIES]), 1.0)
    f_vis2 = 0.0
    if dag == "dag16":
        st2 = rig.download(("stats2",))["stats2"].reshape(-1).view(np.uint64)
        f_vis2 = float(st2[abi.STAT_VISIBLE]) / max(float(st2[abi.STAT_ENTITIES]), 1.0)
    b3 = synth.dag_bytes_per_entity_step(dag, f_vis, f_vis2)
    alg_bytes = b3 * W * E
    achieved = alg_bytes / (kernel_ms * 1e-3) / 1e9 if kernel_ms > 0 else 0.0
    achieved_step = alg_bytes / (step_ms * 1e-3) / 1e9 if step_ms > 0 else 0.0
    traffic, traffic_src = ncu_traffic(dag, W, E, D)
    return {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4),
            "achieved_step": round(achieved_step, 1), "frac_step": round(achieved_step / peak, 4),
            "traffic": traffic, "kernel": rig.executor_text(),
            "algorithmic_bytes_per_launch": int(alg_bytes), "bytes_per_entity_step": round(b3, 3), "f_vis": round(f_vis, 4),
            "kernel_ms": round(kernel_ms, 5), "step_ms": round(step_ms, 5), "peak_source": peak_src,
            "traffic_source": traffic_src,
            "note": "frac: algorithmic bytes / average per-generation device time (CUDA events around each generation); "
                    "frac_step: the same bytes / ms_per_step (gaps between generations included), per GPU. `traffic` = ncu DRAM bytes per "
                    "GENERATION (per launch / generations in the launch). When queued generations share a launch (max_generations_per_launch "
                    "> 1) a block of worlds stays in the 126 MB L2 from one generation to the next, so DRAM traffic falls far below the "
                    "algorithmic bytes and the kernel is bounded by L2 -> SM delivery and instruction issue, not by HBM: frac then says how "
                    "the step compares with streaming its algorithmic bytes at the HBM copy peak (it may exceed 1), and "
                    "one_generation_per_launch is the HBM-bound figure",
            "max_generations_per_launch": int(info.max_generations_per_launch)}


def measure_config(lib, np, torch, abi, synth, dag, W, E, D, K, WU, windows, peak, peak_src, gate: bool) -> dict:
    """A short device-resident measurement of another BASELINE config on this GPU (N = 1), parity-gated the same way."""
    data = synth.generate(W, E, D, dag=dag)
    out = {"workload": f"{W} worlds x {E} entities, depth {D}, {dag}", "steps": K, "warmup": WU, "windows": windows}
    if gate:
        out["parity"] = parity_gate(lib, data, W, E, D, dag, synth.RigOptions(device=torch.cuda.current_device()))
        if not out["parity"]["ok"]:
            return out
    rig = synth.WorldStepRig(lib, W, E, D, data=data, options=synth.RigOptions(device=torch.cuda.current_device()), dag=dag)
    try:
        info = rig.executor_info()
        rig.step(WU)
        spans, kernels, _ = timed_windows(rig, K, windows, torch.cuda.synchronize)
        t_ms = statistics.median(spans)
        kernel_ms = statistics.median(kernels)
        out.update({"value": W * K / (t_ms * 1e-3), "unit": UNIT, "ms_per_step": t_ms / K,
                    "gpu_launches": int(statistics.median(timed_windows.launches)),
                    "generations_per_launch_max": int(info.max_generations_per_launch),
                    "roofline": roofline_of(rig, np, abi, synth, dag, W, E, D, kernel_ms, t_ms / K, peak, peak_src, info)})
    finally:
        rig.close()
    return out


def bind_to_gpu_numa_node(torch, device: int):
    """One process per GPU: run this rank's threads (and so first-touch its pinned buffers) on the CPUs of the NUMA node the GPU's
    PCIe root hangs off.  With several ranks reading back megabytes per step, host buffers on the far socket put every D2H on the
    inter-socket link.  Returns what was done, for the bench line; None when the topology is unknown or the node has no allowed CPU."""
    try:
        pr = torch.cuda.get_device_properties(device)
        bdf = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read().strip())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = cpus & os.sched_getaffinity(0)
        if not allowed or os.environ.get("FIMO_B200_NO_NUMA_BIND"):
            return {"node": node, "bound": False, "cpus_allowed_on_node": len(allowed)}
        os.sched_setaffinity(0, allowed)
        return {"node": node, "bound": True, "cpus": len(allowed)}
    except Exception:
        return None


# ------------------------------------------------------------------------------------------------
# the CUDA arm
# ------------------------------------------------------------------------------------------------
def run_b200(args) -> None:
    import numpy as np
    import torch
    import torch.distributed as dist

    from fimo_b200 import abi, library, sharding, synth

    rank = int(os.environ.get("RANK", "0"))
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the B200 path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    numa = bind_to_gpu_numa_node(torch, local_rank) if world_size > 1 else None  # before any library thread or pinned buffer exists
    distributed = world_size > 1
    if distributed:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    lib = library.load()
    W, E, D = args.worlds, args.entities, args.depth
    K, WU, R = args.steps, max(args.warmup, 3), max(args.windows, 1)
    data = synth.generate(W, E, D, world0=rank * W, dag=args.dag)  # shard = contiguous block of global world indices
    options = synth.RigOptions(fused=not args.no_fusion, graph=not args.no_graph, device=local_rank)
    device = torch.device("cuda", local_rank)

    def barrier():
        if distributed:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce_max(values):
        if not distributed:
            return values
        t = torch.tensor(values, dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(v) for v in t.cpu()]

    # ---- parity gate: this workload, this launch plan and launch shape, against the oracle (rank 0's shard) ----------
    parity = None
    if not args.no_parity_gate:
        if rank == 0:
            parity = parity_gate(lib, data, W, E, D, args.dag, options)
        if distributed:
            flag = torch.tensor([1 if (parity is None or parity["ok"]) else 0], device="cuda")
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            failed = int(flag.item()) == 0
        else:
            failed = not parity["ok"]
        if failed:
            if rank == 0:
                print("bench.py: PARITY GATE FAILED — no bench line:\n" + "\n".join(parity["problems"]), file=sys.stderr, flush=True)
            raise SystemExit(3)

    rig = synth.WorldStepRig(lib, W, E, D, data=data, options=options, dag=args.dag)
    stats_mode = None
    if distributed:  # the statistics exchange is the product's: bound once, before the first generation
        stats_mode = abi.STATS_MODE_PEERS if args.stats == "peers" else abi.STATS_MODE_NCCL
        if stats_mode == abi.STATS_MODE_PEERS:
            sharding.bind_stats_peers(lib, rig.group.handle, dist, torch, slots=64, device=device)
        else:
            sharding.bind_stats_nccl(lib, rig.group.handle, dist, torch, bucket=max(1, args.stats_bucket), device=device)
            WU = max(WU, 2 * max(1, args.stats_bucket))  # warm at least one full bucket's collective
    info = rig.executor_info()

    # ---- value: inputs resident in HBM, windows of K generations back to back ---------------------------------------
    rig.step(WU)
    sampler = ClockSampler(local_rank)
    time.sleep(0.25)
    t0 = time.time()
    spans, kernels, walls = timed_windows(rig, K, R, barrier, reduce_max, sampler)
    window_launches = list(timed_windows.launches)
    t1 = time.time()
    clocks = sampler.summary(t0, t1)
    t_ms = statistics.median(spans)
    kernel_ms = statistics.median(kernels)
    value = world_size * W * K / (t_ms * 1e-3)
    generations_done = WU + K * R

    # ---- N > 1: the exchanged statistics are checked, not just printed --------------------------------------------
    global_stats = None
    if distributed:
        local = rig.stats()
        last = generations_done - 1
        got = sharding.global_stats(lib, rig.group.handle, last, stats_mode)
        want = sharding.allreduce_stats(dist, torch, local, device=device)  # the harness's own NCCL all-reduce of the same words
        if not np.array_equal(got[:5], want[:5]) or int(got[abi.STAT_ENTITIES]) != world_size * W * E or int(got[abi.STAT_WORLDS]) != world_size * W:
            raise SystemExit(f"bench.py: rank {rank}: exchanged statistics {got[:5]} differ from the all-reduced local ones {want[:5]}")
        global_stats = {"visible": int(got[abi.STAT_VISIBLE]), "entities": int(got[abi.STAT_ENTITIES]), "worlds": int(got[abi.STAT_WORLDS]),
                        "world_steps": int(got[abi.STAT_STEPS]), "generation": last, "transport": args.stats,
                        "checked": "== torch.distributed all_reduce(SUM) of every rank's local statistics; entities / worlds == N x shard"}

    # ---- roofline of the dominant kernel -----------------------------------------------------------------------------
    peak, peak_src = measured_peaks()
    roofline = roofline_of(rig, np, abi, synth, args.dag, W, E, D, kernel_ms, t_ms / K, peak, peak_src, info)
    alg_bytes = roofline["algorithmic_bytes_per_launch"]

    # ---- e2e: through the C ABI with HOST buffers, copies inside the timed region --------------------
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(rig, torch, np, abi, W, E, K, distributed, dist, world_size, roofline["f_vis"])

    default_config = (W, E, D, args.dag) == (4096, 1024, 4, "chain3")
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world_size, "steps": K, "warmup": WU,
        "ms_per_step": t_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": config_of(W, E, D, args.dag),
        "run": {"executor": ("fused world-step kernel" if info.fused_world_step else "one kernel per system")
                            + (" in a CUDA graph" if info.uses_cuda_graph else ", plain stream launches"),
                "state_touched_per_step_mb": round(alg_bytes / 1e6, 1),
                "timing": f"{R} windows of exactly {K} generations, each bracketed by barrier + synchronize, device-timed (CUDA events on the "
                          "group's stream), max over ranks; value = median window",
                "sharding": "worlds by global index, no data-path collective; every generation's statistics (64 B) exchanged by the product: "
                            + ("peer stores over NVLink from the publishing CTA" if args.stats == "peers" else
                               f"ncclAllReduce on a library side stream, {args.stats_bucket} generations per collective")},
        "windows": {"n": R, "ms_per_step_median": t_ms / K, "ms_per_step_min": min(spans) / K, "ms_per_step_max": max(spans) / K,
                    "value_min": world_size * W * K / (max(spans) * 1e-3), "value_max": world_size * W * K / (min(spans) * 1e-3)},
        "wall_ms_per_step": statistics.median(walls) / K, "device_ms_per_step": kernel_ms,
        "gpu_launches": int(statistics.median(window_launches)),
        "gpu_launches_note": f"kernel launches inside one timed window of {K} generations, counted by the executor (per window: {window_launches}); "
                             f"one launch per generation would be {int(info.kernel_launches_per_generation) * K}: generations queued behind two "
                             f"running launches leave together, up to {int(info.max_generations_per_launch)} per launch",
        "roofline": roofline, "clocks": clocks,
    }
    if parity is not None:
        line["parity"] = {k: parity[k] for k in ("ok", "checked", "seconds", "steps", "kernel_launches", "covers_shared_launches")}
    if e2e is not None:
        line["e2e"] = e2e["headline"]
        line["e2e_constants_only"] = e2e["constants_only"]
        line["e2e_full_state"] = e2e["full_state"]
    if global_stats is not None:
        line["global_stats"] = global_stats
    if numa is not None:
        line["run"]["numa"] = numa
    if distributed:
        dist.barrier()  # every rank has finished storing into its peers' inboxes before anyone tears down
    rig.close()
    if rank == 0 and world_size == 1 and default_config and not args.no_other_configs:
        # the other single-GPU BASELINE shapes, short, so that the driver's record carries them (not bench lines of their own)
        line["other_configs"] = {
            "configs[2]": measure_config(lib, np, torch, abi, synth, "dag8", 256, 65536, 8, 20, 5, 3, peak, peak_src, not args.no_parity_gate),
            "configs[3] per-GPU shape": measure_config(lib, np, torch, abi, synth, "chain3", 4096, 4096, 4, 20, 5, 3, peak, peak_src, not args.no_parity_gate),
            "configs[4] per-GPU shape": measure_config(lib, np, torch, abi, synth, "dag16", 1024, 1024, 4, 20, 5, 3, peak, peak_src, not args.no_parity_gate),
        }
        # the same workload with ONE generation per launch (FIMO_B200_GENS=1, read when the plan is made): every generation then
        # streams its whole state through HBM, which is what `bound: hbm` describes; the default lets queued generations share a launch
        os.environ["FIMO_B200_GENS"] = "1"
        try:
            one = measure_config(lib, np, torch, abi, synth, args.dag, W, E, D, K, WU, 3, peak, peak_src, False)
        finally:
            del os.environ["FIMO_B200_GENS"]
        line["roofline"]["one_generation_per_launch"] = {
            "value": one["value"], "ms_per_step": one["ms_per_step"], "gpu_launches": one["gpu_launches"],
            "achieved": one["roofline"]["achieved"], "frac": one["roofline"]["frac"], "frac_step": one["roofline"]["frac_step"]}
        for name, cfg in line["other_configs"].items():
            if "parity" in cfg and not cfg["parity"]["ok"]:
                print(f"bench.py: PARITY GATE FAILED for {name} — no bench line:\n" + "\n".join(cfg["parity"]["problems"]), file=sys.stderr, flush=True)
                raise SystemExit(3)
        # the same figures as top-level scalars (a record that keeps only scalars still shows every BASELINE shape)
        line["c2_one_gen_frac"], line["c2_one_gen_value"] = one["roofline"]["frac"], one["value"]
        for key, name in (("c3", "configs[2]"), ("c4shape", "configs[3] per-GPU shape"), ("c5shape", "configs[4] per-GPU shape")):
            cfg = line["other_configs"][name]
            line[key + "_value"], line[key + "_frac"], line[key + "_frac_step"] = cfg["value"], cfg["roofline"]["frac"], cfg["roofline"]["frac_step"]
            line[key + "_parity_ok"] = cfg["parity"]["ok"] if "parity" in cfg else None
    if rank == 0 and not args.no_cpu_baseline and world_size == 1:
        base = cpu_arm(max(1, min(CPU_SAMPLE_WORLDS, W, (CPU_SAMPLE_WORLDS * 1024) // max(E, 1))), E, D, args.cpu_steps, 2, args.dag, seconds_budget=10.0)
        line["cpu_baseline"] = {k: base[k] for k in ("value", "unit", "cores", "kind", "sample")}
        if default_config:  # BASELINE.json configs[0]: a single world x 1k entities on the CPU workers
            c1 = cpu_arm(1, 1024, 4, 0, 20, "chain3", seconds_budget=3.0)
            line["cpu_baseline_configs0"] = {k: c1[k] for k in ("value", "unit", "cores", "kind", "sample")}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if distributed:
        dist.barrier()
        dist.destroy_process_group()


def run_e2e(rig, torch, np, abi, W, E, K, distributed, dist, world_size, f_vis) -> dict:
    """The same metric end to end: every step copies that step's inputs host->device from pinned memory and reads the
    step's result device->host, all through the C ABI, several steps in flight (schedule is asynchronous in the
    reference too, SystemGroup.zig:430-452), every result waited for inside the timed region.

    headline (`e2e`): inputs = camera frustum + dt of every world (what a host changes each frame; the entity state
                stays resident like model weights); result = WHAT CULL PRODUCES: per-world visible counts, the
                visibility masks and the ordered visible lists packed back to back (fimo_b200_group_schedule_readback).
    constants_only: same inputs, result = statistics + per-world counts only (fimo_b200_group_schedule_io).
    full_state: every live-in resource uploaded and every live-out resource downloaded each step — the worst case.
    Each leg is warmed for >= 2 x the pipeline depth, then timed for at least 200 steps and 0.4 s, three times; the
    median is reported."""
    from fimo_b200 import synth
    from fimo_b200.worlds import Fence

    def pinned_like(a):
        t = torch.empty(a.shape, dtype=torch.float32 if a.dtype == np.float32 else torch.int32).pin_memory()
        v = t.numpy().view(a.dtype)
        v[...] = a
        return t, v

    keep = []
    DEPTH = max(1, int(os.environ.get("FIMO_B200_E2E_DEPTH", "4")))  # steps in flight (A/B: more depth hides host jitter, costs latency)
    in_slots, out_slots, rb_slots, fences = [], [], [], [Fence(rig.lib) for _ in range(DEPTH)]
    capacity = int(1.15 * f_vis * W * E) + 1024  # list entries the read-back buffer holds (measured visible fraction + 15 %)
    rb_words = rig.readback_words(capacity)
    for _ in range(DEPTH):
        si, so = {}, {}
        for name in ("frustum", "dt"):
            t, v = pinned_like(rig.host[name])
            keep.append(t)
            si[name] = v
        for name in ("stats", "vis_count"):
            t, v = pinned_like(rig.host[name])
            keep.append(t)
            so[name] = v
        t = torch.zeros(rb_words, dtype=torch.int32).pin_memory()
        keep.append(t)
        in_slots.append(si), out_slots.append(so), rb_slots.append(t.numpy().view(np.uint32))
    up_lists = [rig.io_list(in_slots[i]) for i in range(DEPTH)]
    down_lists = [rig.io_list(out_slots[i]) for i in range(DEPTH)]
    rb_descs = [rig.readback_desc(rb_slots[i]) for i in range(DEPTH)]

    def make_pipeline(submit):
        state = {"i": 0, "pending": [False] * DEPTH}

        def step():
            slot = state["i"] % DEPTH
            state["i"] += 1
            if state["pending"][slot]:
                fences[slot].wait()  # the result of the step that used this slot DEPTH steps ago has landed
            fences[slot].reset()
            submit(slot)
            state["pending"][slot] = True

        def drain():
            for slot in range(DEPTH):
                if state["pending"][slot]:
                    fences[slot].wait()
                    state["pending"][slot] = False

        return step, drain

    def timed_leg(step, drain, min_steps=200, min_seconds=0.4, repeats=3):
        for _ in range(3 * DEPTH):
            step()
        drain()
        rates = []
        for _ in range(repeats):
            if distributed:
                dist.barrier()
            torch.cuda.synchronize()
            n, t0 = 0, time.perf_counter()
            while n < min_steps or time.perf_counter() - t0 < min_seconds:
                step()
                n += 1
            drain()
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            if distributed:  # all ranks run the same number of steps per second of the slowest: rate = min over ranks
                tt = torch.tensor([n / dt], dtype=torch.float64, device="cuda")
                dist.all_reduce(tt, op=dist.ReduceOp.MIN)
                rates.append(float(tt.item()))
            else:
                rates.append(n / dt)
        return statistics.median(rates), min(rates), max(rates), n

    # headline: cull outputs come back
    step_rb, drain_rb = make_pipeline(lambda s: rig.schedule_readback(up_lists[s], rb_descs[s], fences[s]))
    rate, lo, hi, n = timed_leg(step_rb, drain_rb)
    last = rig.parse_readback(rb_slots[0])
    if last["truncated"] or last["n_worlds"] != W or int(last["counts"].sum()) != last["total"]:
        raise SystemExit(f"bench.py: read-back buffer inconsistent: {last['total']} visible, truncated={last['truncated']}")
    h2d = int(sum(v.nbytes for v in in_slots[0].values()))
    head = {"value": world_size * W * rate, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": int(rb_words * 4),
            "ms_per_step": 1e3 / rate, "value_min": world_size * W * lo, "value_max": world_size * W * hi, "steps_per_repeat": n, "repeats": 3,
            "inputs": "frustum + dt of every world from pinned host memory",
            "result": f"what cull produces: per-world visible counts + offsets, visibility masks, ordered visible lists packed back to back "
                      f"({last['total']} entries of a {capacity}-entry buffer)",
            "api": "fimo_b200_group_schedule_readback: one C-ABI call per step = H2D of the step's inputs (staged beside the running generation) "
                   "+ one generation (the fused kernel leaves a second copy of its cull outputs) + scan-and-pack of that copy and ONE copy-engine D2H, both on the read-back stream beside the next generation + completion fence; "
                   f"{DEPTH} steps in flight, every result waited for inside the timed region",
            "in_flight": DEPTH}

    # constants only (round 1's headline)
    step_io, drain_io = make_pipeline(lambda s: rig.schedule_io(up_lists[s], down_lists[s], fences[s]))
    rate_c, lo_c, hi_c, n_c = timed_leg(step_io, drain_io)

    def step_io_sync():
        fences[0].reset()
        rig.schedule_io(up_lists[0], down_lists[0], fences[0])
        fences[0].wait()

    rate_s, _, _, _ = timed_leg(step_io_sync, lambda: None, min_steps=100, min_seconds=0.2)
    consts = {"value": world_size * W * rate_c, "unit": UNIT, "h2d_bytes_per_step": h2d,
              "d2h_bytes_per_step": int(sum(v.nbytes for v in out_slots[0].values())), "ms_per_step": 1e3 / rate_c,
              "value_min": world_size * W * lo_c, "value_max": world_size * W * hi_c, "steps_per_repeat": n_c,
              "result": "statistics + per-world visible counts", "api": f"fimo_b200_group_schedule_io, {DEPTH} steps in flight",
              "sync_value": world_size * W * rate_s, "sync_api": "same call, one step in flight: the host waits for each result before queuing the next"}

    # full state
    fins, fouts = {}, {}
    for name in synth.LIVE_IN + ("frustum", "dt"):
        t, v = pinned_like(rig.host[name])
        keep.append(t)
        fins[name] = v
    for name in synth.LIVE_OUT + ("vis_count", "stats"):
        t, v = pinned_like(rig.host[name])
        keep.append(t)
        fouts[name] = v

    def step_full():
        for name, v in fins.items():
            rig.upload(name, v)
        rig.step(1)
        for name, v in fouts.items():
            rig.download_into(name, v)

    step_full()
    torch.cuda.synchronize()
    fsteps = 4
    t0 = time.perf_counter()
    for _ in range(fsteps):
        step_full()
    torch.cuda.synchronize()
    dt_full = time.perf_counter() - t0
    full = {"value": world_size * W * fsteps / dt_full, "unit": UNIT,
            "h2d_bytes_per_step": int(sum(v.nbytes for v in fins.values())),
            "d2h_bytes_per_step": int(sum(v.nbytes for v in fouts.values())), "steps": fsteps,
            "ms_per_step": 1e3 * dt_full / fsteps}
    return {"headline": head, "constants_only": consts, "full_state": full}


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
