#!/bin/bash
# 2 GPUs, final tree: the driver's bench command with the shadow read-back e2e leg; multi-GPU parity tests.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_multi_gpu.py -m gpu -x -q --timeout 120 > gpurun_out/r3c_pytest_2gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r3c_pytest_2gpu.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r3c_bench_c2_2gpu.json 2> gpurun_out/r3c_bench_c2_2gpu.err; echo "2gpu rc=$?"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 bench.py --impl reference --gpus 2 --steps 20 --warmup 5 > gpurun_out/r3c_bench_c2_2gpu_reference.json 2>/dev/null; echo "2gpu ref rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r3c_bench_c2_2gpu.json").read().strip().splitlines()[-1])
print("N", d["n_gpus"], round(d["value"]/1e6,2), d["roofline"]["frac"], "e2e", round(d["e2e"]["value"]/1e6,2), "const", round(d["e2e_constants_only"]["value"]/1e6,2), d["clocks"], d["global_stats"]["transport"])
r = json.loads(open("gpurun_out/r3c_bench_c2_2gpu_reference.json").read().strip().splitlines()[-1])
print("reference", r.get("value"), r.get("n_gpus"), r.get("impl"))
PY
