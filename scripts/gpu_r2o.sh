#!/bin/bash
# Team mode at 5 CTAs/SM (80-register build, 2-deep ring) against the default 4 CTAs/SM x 3-deep ring, same box.
mkdir -p gpurun_out
S="--steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-other-configs"
C3="--dag dag8 --worlds 256 --entities 65536 --depth 8"
timeout 120 python bench.py $S $C3 --no-parity-gate > gpurun_out/r2o_bench_c3_default.json 2>/dev/null; echo "default rc=$?"
FIMO_B200_PIPE=128,5,2 timeout 120 python bench.py $S $C3 > gpurun_out/r2o_bench_c3_5x2.json 2> gpurun_out/r2o_5x2.err; echo "5x2 rc=$?"; tail -2 gpurun_out/r2o_5x2.err
FIMO_B200_PIPE=128,4,2 timeout 120 python bench.py $S $C3 --no-parity-gate > gpurun_out/r2o_bench_c3_4x2.json 2>/dev/null; echo "4x2 rc=$?"
FIMO_B200_PIPE=128,5,2 FIMO_B200_TEAM=16 timeout 120 python bench.py $S $C3 --no-parity-gate > gpurun_out/r2o_bench_c3_5x2_team16.json 2>/dev/null; echo "5x2 t16 rc=$?"
FIMO_B200_PIPE=128,5,2 FIMO_B200_TEAM=12 timeout 120 python bench.py $S $C3 --no-parity-gate > gpurun_out/r2o_bench_c3_5x2_team12.json 2>/dev/null; echo "5x2 t12 rc=$?"
FIMO_B200_PIPE=128,5,2 timeout 120 python bench.py $S --worlds 256 --entities 65536 --depth 8 > gpurun_out/r2o_bench_c3_chain3_5x2.json 2>/dev/null; echo "chain3 5x2 rc=$?"
FIMO_B200_PIPE=128,6,2 timeout 120 python bench.py $S --worlds 256 --entities 65536 --depth 8 > gpurun_out/r2o_bench_c3_chain3_6x2.json 2>/dev/null; echo "chain3 6x2 rc=$?"
FIMO_B200_PIPE=128,5,2 timeout 120 python bench.py $S --dag dag8 --worlds 592 --entities 65536 --depth 8 --no-parity-gate > gpurun_out/r2o_bench_592x64k_5x2.json 2>/dev/null; echo "592 rc=$?"
for f in gpurun_out/r2o_bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r = d.get("roofline", {})
    print(sys.argv[1].split("/")[-1], round(d["value"] / 1e6, 3), "M", "frac", r.get("frac"), "parity", (d.get("parity") or {}).get("ok"), r.get("kernel", "")[:110])
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
