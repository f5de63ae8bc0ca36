#!/bin/bash
# e2e legs with 4 (default), 6, 8 and 2 steps in flight.
mkdir -p gpurun_out
for d in 4 8 6 2; do FIMO_B200_E2E_DEPTH=$d timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-other-configs --no-parity-gate > gpurun_out/r2x_bench_c2_depth$d.json 2>/dev/null; echo "depth $d rc=$?"; done
for f in gpurun_out/r2x_bench_*.json; do python - "$f" <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1].split("/")[-1], "e2e", round(d["e2e"]["value"] / 1e6, 2), "M", "const-only", round(d["e2e_constants_only"]["value"] / 1e6, 2), "sync", round(d["e2e_constants_only"]["sync_value"] / 1e6, 2))
PY
done
