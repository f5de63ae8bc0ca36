#!/bin/bash
# 8 GPUs, final tree: the driver's bench command (value + shadow read-back e2e).
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29528 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r3d_bench_c2_8gpu.json 2> gpurun_out/r3d_bench_c2_8gpu.err; echo "8gpu rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r3d_bench_c2_8gpu.json").read().strip().splitlines()[-1])
print("N", d["n_gpus"], round(d["value"]/1e6,2), d["roofline"]["frac"], d["roofline"]["frac_step"], "e2e", round(d["e2e"]["value"]/1e6,2), "const", round(d["e2e_constants_only"]["value"]/1e6,2), d["clocks"], d["config"].get("numa"))
PY
