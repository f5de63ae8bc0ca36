#!/bin/bash
# Round-2, 2 GPUs: N>1 parity (two shards == one batch on the oracle, exchanged statistics == oracle's), the driver's bench
# command at N=2 for both statistics transports, and the configs[4] shape.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q --timeout 120 > gpurun_out/r2z_pytest_2gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2z_pytest_2gpu.log
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5"
timeout 300 $T > gpurun_out/r2z_bench_c2_2gpu.json 2> gpurun_out/r2z_bench_c2_2gpu.err; echo "2gpu peers rc=$?"
timeout 300 $T --stats nccl --no-e2e > gpurun_out/r2z_bench_c2_2gpu_nccl.json 2> gpurun_out/r2z_bench_c2_2gpu_nccl.err; echo "2gpu nccl rc=$?"
timeout 300 $T --dag dag16 --worlds 1024 --no-e2e > gpurun_out/r2z_bench_c5_2gpu.json 2> gpurun_out/r2z_bench_c5_2gpu.err; echo "2gpu c5 rc=$?"
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-other-configs > gpurun_out/r2z_bench_c2_1gpu.json 2>/dev/null; echo "1gpu rc=$?"
for f in gpurun_out/r2z_bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r = d.get("roofline", {})
    print(sys.argv[1].split("/")[-1], "N", d["n_gpus"], round(d["value"] / 1e6, 3), "M", "frac", r.get("frac"), "frac_step", r.get("frac_step"), "e2e", (d.get("e2e") or {}).get("value"),
          "launches", d.get("gpu_launches"), "stats", (d.get("global_stats") or {}).get("transport"))
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
