#!/bin/bash
# end-of-round verification: full GPU suite, smoke, and the bench lines kept under profiles/
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/final_c2.json 2> gpurun_out/final_c2.err
timeout 300 python bench.py --no-cpu-baseline --dag dag8 --worlds 256 --entities 65536 --depth 8 > gpurun_out/final_c3_dag8.json 2>> gpurun_out/final.err
timeout 300 python bench.py --no-cpu-baseline --dag dag16 --worlds 1024 --entities 1024 > gpurun_out/final_c5_dag16.json 2>> gpurun_out/final.err
timeout 300 python bench.py --no-cpu-baseline --dag dag8 > gpurun_out/final_dag8_4096.json 2>> gpurun_out/final.err
python - <<'P'
import json
for n in ("c2","c3_dag8","c5_dag16","dag8_4096"):
    try:
        d=json.loads(open(f"gpurun_out/final_{n}.json").read().strip().splitlines()[-1])
        print(n, d["value"], d["roofline"]["frac"], d["e2e"]["value"], (d.get("cpu_baseline") or {}).get("value"), d["gpu_launches"])
    except Exception as e: print(n, "failed", e)
P
