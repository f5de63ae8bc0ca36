#!/bin/bash
# Team mode: parent rows of the next chunk fetched by the TMA engine (tile::gather4) instead of cp.async through the LSU.
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q --timeout 60 -k "team or full_size or large_world or dags or empty" > gpurun_out/r2u_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2u_pytest.log
S="--steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-other-configs"
C3="--dag dag8 --worlds 256 --entities 65536 --depth 8"
timeout 100 python bench.py $S $C3 > gpurun_out/r2u_bench_c3.json 2>/dev/null; echo "c3 rc=$?"
FIMO_B200_NO_ROW_GATHER=1 timeout 100 python bench.py $S $C3 --no-parity-gate > gpurun_out/r2u_bench_c3_cpasync.json 2>/dev/null; echo "c3 cp.async rc=$?"
timeout 100 python bench.py $S --worlds 256 --entities 65536 --depth 8 > gpurun_out/r2u_bench_c3_chain3.json 2>/dev/null; echo "c3 chain3 rc=$?"
timeout 100 python bench.py $S --dag dag8 --worlds 592 --entities 65536 --depth 8 --no-parity-gate > gpurun_out/r2u_bench_592x64k.json 2>/dev/null; echo "592 rc=$?"
timeout 100 python bench.py $S --dag dag8 --worlds 256 --entities 65536 --depth 2 --no-parity-gate > gpurun_out/r2u_bench_c3_depth2.json 2>/dev/null; echo "d2 rc=$?"
for f in gpurun_out/r2u_bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1].split("/")[-1], round(d["value"] / 1e6, 4), "M", "frac", d["roofline"]["frac"], "parity", (d.get("parity") or {}).get("ok"))
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
