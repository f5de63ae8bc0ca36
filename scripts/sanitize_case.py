#!/usr/bin/env python
"""One small world-step run for compute-sanitizer (memcheck / racecheck / synccheck), checked against the oracle:
   compute-sanitizer --tool racecheck python scripts/sanitize_case.py chain3 16 1024 4 2 [VAR=value ...]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
dag, W, E, D, K = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
for kv in sys.argv[6:]:
    k, v = kv.split("=", 1)
    os.environ[k] = v

from fimo_b200 import library, synth  # noqa: E402
from oracle import loader  # noqa: E402

data = synth.generate(W, E, D, dag=dag)
rig = synth.WorldStepRig(library.load(), W, E, D, data={k: v.copy() for k, v in data.items()}, dag=dag)
info = rig.executor_info()
rig.step(K)
got = rig.download()
rig.close()
want = synth.step_in_slices(loader.load(), data, E, D, K, dag=dag, slices=1, workers=2)
problems = synth.compare_outputs(got, want, 0.0)
print(f"{dag} {W}x{E} depth {D} K={K}: fused={info.fused_world_step} kernels/generation={info.kernel_launches_per_generation} "
      f"parity={'ok' if not problems else problems}")
sys.exit(1 if problems else 0)
