#!/bin/bash
mkdir -p gpurun_out
SECONDS=0; timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2y_bench_c2.json 2> gpurun_out/r2y_bench_c2.err; echo "bench rc=$?"; echo "wall ${SECONDS}s"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2y_bench_c2.json").read().strip().splitlines()[-1])
print(round(d["value"]/1e6,2), d["roofline"]["frac"], round(d["e2e"]["value"]/1e6,2), d["gpu_launches"], d["clocks"])
for k, v in d["other_configs"].items(): print(k, round(v["value"]/1e6,3), v["roofline"]["frac"], v["parity"]["ok"], v["parity"]["seconds"])
print(d["cpu_baseline"]["value"], d["cpu_baseline_configs0"]["value"])
PY
