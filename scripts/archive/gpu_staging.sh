#!/bin/bash
# staged uploads: full GPU suite, then the default bench line with and without staging
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu 2>&1 | tail -5 > gpurun_out/staging_tests.log
timeout 300 python bench.py > gpurun_out/staging_on.json 2> gpurun_out/staging_on.err
FIMO_B200_IO_NO_STAGING=1 timeout 300 python bench.py > gpurun_out/staging_off.json 2> gpurun_out/staging_off.err
cat gpurun_out/staging_tests.log
python - <<'P'
import json
for n in ("on","off"):
    try:
        d=json.loads(open(f"gpurun_out/staging_{n}.json").read().strip().splitlines()[-1])
        print(n, d["value"], d["roofline"]["frac"], d["e2e"]["value"], d["e2e"]["sync_value"], d["cpu_baseline"]["value"])
    except Exception as e: print(n, "failed", e)
P
