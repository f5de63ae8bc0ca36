#!/bin/bash
# Round-2 call 7: full GPU suite on the current tree, the driver's bench line, the reference arm, other shapes, profiles.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2g_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2g_pytest.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r2g_bench_c2.json 2> gpurun_out/r2g_bench_c2.err; echo "bench rc=$?"
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2g_bench_c2_reference_arm.json 2>/dev/null; echo "ref rc=$?"
S="--steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-other-configs"
python bench.py $S --entities 4096 > gpurun_out/r2g_bench_c4_shape.json 2>/dev/null; echo "c4 rc=$?"
python bench.py $S --dag dag16 --worlds 1024 > gpurun_out/r2g_bench_c5_shape.json 2>/dev/null; echo "c5 rc=$?"
python bench.py $S --dag dag8 --worlds 256 --entities 65536 --depth 8 > gpurun_out/r2g_bench_c3.json 2>/dev/null; echo "c3 rc=$?"
python bench.py $S --dag dag8 > gpurun_out/r2g_bench_dag8_4096.json 2>/dev/null; echo "dag8 rc=$?"
python bench.py $S --dag dag16 > gpurun_out/r2g_bench_dag16_4096.json 2>/dev/null; echo "dag16 rc=$?"
for f in gpurun_out/r2g_bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r = d.get("roofline", {})
    print(sys.argv[1].split("/")[-1], round(d["value"] / 1e6, 2), "M", "frac", r.get("frac"), "frac_step", r.get("frac_step"), "e2e", (d.get("e2e") or {}).get("value"), "launches", d.get("gpu_launches"))
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
bash scripts/gpu_r2_profile.sh
