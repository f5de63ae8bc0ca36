#!/bin/bash
# Round-2 call 9: set 2 compiled at 80 registers (no spilled prev_copy loads); the driver's bench line with the
# one-generation-per-launch leg and the repaired CPU arm; the reference arm.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2i_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2i_pytest.log
S="--steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-other-configs"
python bench.py $S --dag dag16 --worlds 1024 > gpurun_out/r2i_bench_c5_shape.json 2>/dev/null; echo "c5 rc=$?"
python bench.py $S --dag dag16 > gpurun_out/r2i_bench_dag16_4096.json 2>/dev/null; echo "dag16 rc=$?"
python bench.py --steps 20 --warmup 5 > gpurun_out/r2i_bench_c2.json 2> gpurun_out/r2i_bench_c2.err; echo "bench rc=$?"
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2i_bench_c2_reference_arm.json 2>/dev/null; echo "ref rc=$?"
for f in gpurun_out/r2i_bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r = d.get("roofline", {})
    print(sys.argv[1].split("/")[-1], round(d["value"] / 1e6, 3), "M", "frac", r.get("frac"), "frac_step", r.get("frac_step"), "e2e", (d.get("e2e") or {}).get("value"),
          "launches", d.get("gpu_launches"), "one-gen", (r.get("one_generation_per_launch") or {}).get("frac"), "cpu", (d.get("cpu_baseline") or {}).get("value"))
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
