#!/bin/bash
# Round-2 call: team mode with the parent-plane look-ahead ring (C3), prefetch-mode A/B now that rows are L2-resident.
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -x -q --timeout 100 --timeout-method=thread > gpurun_out/r2l_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2l_pytest.log
S="--steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-other-configs"
C3="--dag dag8 --worlds 256 --entities 65536 --depth 8"
timeout 120 python bench.py $S $C3 > gpurun_out/r2l_bench_c3.json 2>/dev/null; echo "c3 rc=$?"
FIMO_B200_PREFETCH=1 timeout 120 python bench.py $S $C3 --no-parity-gate > gpurun_out/r2l_bench_c3_prefetch1.json 2>/dev/null; echo "c3 p1 rc=$?"
FIMO_B200_PREFETCH=0 timeout 120 python bench.py $S $C3 --no-parity-gate > gpurun_out/r2l_bench_c3_prefetch0.json 2>/dev/null; echo "c3 p0 rc=$?"
FIMO_B200_PUBLISH=chunk timeout 120 python bench.py $S $C3 --no-parity-gate > gpurun_out/r2l_bench_c3_publish_chunk.json 2>/dev/null; echo "c3 pc rc=$?"
timeout 120 python bench.py $S --worlds 256 --entities 65536 --depth 8 > gpurun_out/r2l_bench_c3_chain3.json 2>/dev/null; echo "c3 chain3 rc=$?"
timeout 120 python bench.py $S --dag dag8 --worlds 592 --entities 65536 --depth 8 --no-parity-gate > gpurun_out/r2l_bench_592x64k.json 2>/dev/null; echo "592 rc=$?"
for f in gpurun_out/r2l_bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r = d.get("roofline", {})
    print(sys.argv[1].split("/")[-1], round(d["value"] / 1e6, 3), "M", "frac", r.get("frac"), "frac_step", r.get("frac_step"), "launches", d.get("gpu_launches"), "parity", (d.get("parity") or {}).get("ok"))
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
