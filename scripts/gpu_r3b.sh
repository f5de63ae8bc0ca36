#!/bin/bash
# Read-back with the pack beside the next generation (shadow copy of the cull outputs) against the in-stream pack.
mkdir -p gpurun_out
timeout 500 python -m pytest tests -m gpu -x -q --timeout 120 > gpurun_out/r3b_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r3b_pytest.log
B="timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-other-configs"
$B > gpurun_out/r3b_bench_c2.json 2> gpurun_out/r3b_bench_c2.err; echo "shadow rc=$?"
FIMO_B200_IO_COPY_KERNEL=1 $B > gpurun_out/r3b_bench_c2_copykernel.json 2>/dev/null; echo "in-stream rc=$?"
$B > gpurun_out/r3b_bench_c2_again.json 2>/dev/null; echo "shadow again rc=$?"
for f in gpurun_out/r3b_bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1].split("/")[-1], "value", round(d["value"] / 1e6, 2), "frac", d["roofline"]["frac"], "one-gen", (d["roofline"].get("one_generation_per_launch") or {}).get("frac"), "e2e", round(d["e2e"]["value"] / 1e6, 2), "const", round(d["e2e_constants_only"]["value"] / 1e6, 2))
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
