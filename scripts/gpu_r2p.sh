#!/bin/bash
# Set 2 without the spilled prev_copy load (mask-word sums accumulated in shared memory): configs[4] shape and dag16 4096 x 1024.
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -x -q -k "dag16 or full_size or queued or dags or stats_ring" > gpurun_out/r2p_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2p_pytest.log
S="--steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-other-configs"
timeout 120 python bench.py $S --dag dag16 --worlds 1024 > gpurun_out/r2p_bench_c5_shape.json 2>/dev/null; echo "c5 rc=$?"
timeout 120 python bench.py $S --dag dag16 > gpurun_out/r2p_bench_dag16_4096.json 2>/dev/null; echo "dag16 rc=$?"
FIMO_B200_GENS=1 timeout 120 python bench.py $S --dag dag16 --worlds 1024 --no-parity-gate > gpurun_out/r2p_bench_c5_shape_onegen.json 2>/dev/null; echo "c5 1gen rc=$?"
for f in gpurun_out/r2p_bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r = d.get("roofline", {})
    print(sys.argv[1].split("/")[-1], round(d["value"] / 1e6, 3), "M", "frac", r.get("frac"), "frac_step", r.get("frac_step"), "launches", d.get("gpu_launches"), "parity", (d.get("parity") or {}).get("ok"))
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
