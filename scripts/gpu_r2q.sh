#!/bin/bash
# C3: pause between polls of a level counter (hand-off latency).
mkdir -p gpurun_out
S="--steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-other-configs --no-parity-gate"
C3="--dag dag8 --worlds 256 --entities 65536 --depth 8"
for ns in 64 0 16 256; do FIMO_B200_POLL_NS=$ns timeout 120 python bench.py $S $C3 > gpurun_out/r2q_bench_c3_poll$ns.json 2>/dev/null; echo "poll $ns rc=$?"; done
for f in gpurun_out/r2q_bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1].split("/")[-1], round(d["value"] / 1e6, 4), "M", "frac", d["roofline"]["frac"])
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
