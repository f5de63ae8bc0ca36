#!/usr/bin/env python
"""bench.py — batched world-steps/sec of the fimo_worlds system-group step on B200 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W]            the CUDA product, one process per GPU
  python bench.py --impl reference [...]                         the reference's CPU path (oracle restatement) on host cores

A "step" is one generation (`system_group_schedule`) of the 3-system chain integrate -> transform propagate
-> AABB frustum cull over one batch of synthetic worlds.  N=1 workload = BASELINE.json configs[1]:
4096 worlds x 1024 entities, depth-4 hierarchy.  N>1: the same batch per GPU (weak scaling), worlds sharded by
global index, no data-path collective; the end-of-step statistics are all-reduced over NCCL on a side stream.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
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


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=200)
    p.add_argument("--warmup", type=int, default=20)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--worlds", type=int, default=WORLDS_PER_GPU, help="worlds per GPU")
    p.add_argument("--entities", type=int, default=ENTITIES)
    p.add_argument("--depth", type=int, default=DEPTH)
    p.add_argument("--no-fusion", action="store_true", help="one kernel per system (CUDA graph of the DAG)")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--cpu-steps", type=int, default=0, help="steps of the CPU arm (0 = sized for ~15 s)")
    return p.parse_args()


def measured_peaks() -> tuple[float, str]:
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference's own structure — one World + one SystemGroup per world, every group scheduled
# each step on ONE shared pool with all host threads, systems as system_run callbacks.
# ------------------------------------------------------------------------------------------------
def cpu_arm(worlds: int, entities: int, depth: int, steps: int, warmup: int) -> dict:
    from fimo_b200 import synth
    from fimo_b200.worlds import create_pool
    from oracle import loader

    lib = loader.load()
    cores = int(lib.fimo_oracle_hardware_threads()) or (os.cpu_count() or 1)
    pool = create_pool(lib, "cpu-baseline", cores)
    data = synth.generate(worlds, entities, depth)
    rigs = []
    for w in range(worlds):
        one = {k: (v if k == "stats" else v[w:w + 1]).copy() for k, v in data.items()}
        rigs.append(synth.WorldStepRig(lib, 1, entities, depth, data=one, pool=pool))
    groups = (C.c_void_p * worlds)(*[r.group.handle for r in rigs])
    if steps <= 0:  # size the run for ~15 s of CPU work
        t = lib.fimo_oracle_step_groups(groups, worlds, 2)
        steps = max(3, min(2000, int(15.0 / max(t / 2, 1e-6))))
    lib.fimo_oracle_step_groups(groups, worlds, max(warmup, 1))
    seconds = lib.fimo_oracle_step_groups(groups, worlds, steps)
    for r in rigs:
        r.close()
    pool.unref()
    return {
        "value": worlds * steps / seconds, "unit": UNIT, "cores": cores, "kind": "port",
        "sample": f"{worlds} worlds x {entities} entities, depth {depth}, 3-system chain, {steps} steps, one World+SystemGroup "
                  f"per world on one shared {cores}-thread pool (oracle restatement of fimo_worlds/fimo_tasks; the reference "
                  f"itself needs zig + boost.context and cannot be built here)",
        "seconds": seconds, "steps": steps, "ms_per_step": 1e3 * seconds / steps,
    }


def run_reference(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = args.steps if args.steps != 200 else 0
    base = cpu_arm(CPU_SAMPLE_WORLDS, args.entities, args.depth, min(steps, 400), min(args.warmup, 3))
    line = {
        "impl": "reference", "metric": METRIC, "value": base["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": base["steps"], "warmup": min(args.warmup, 3), "ms_per_step": base["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.worlds} worlds x {args.entities} entities, depth {args.depth}, 3-system chain "
                               f"(integrate -> propagate -> cull); CPU arm steps a {CPU_SAMPLE_WORLDS}-world sample per step"},
        "cpu_baseline": {k: base[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# clocks sampled DURING the timed region
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device: int):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(device)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def summary(self, t0: float, t1: float) -> dict:
        if self.proc is not None:
            self.proc.terminate()
        rows = [r for t, r in self.rows if t0 <= t <= t1 + 0.15] or [r for _, r in self.rows[-3:]]
        sm = sorted(int(float(r[1])) for r in rows if len(r) > 2 and r[1].replace(".", "").isdigit())
        mx = [int(float(r[2])) for r in rows if len(r) > 2 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in rows if len(r) >= 9 for n, v in zip(names, r[5:9]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(rows)}
