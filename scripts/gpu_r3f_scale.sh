#!/bin/bash
# Final tree, one 8-GPU box: the driver's scaling sweep (N = 1, 2, 4, 8 back to back, --steps 20 --warmup 5), both arms at N = 1.
mkdir -p gpurun_out
run() { n=$1
    if [ "$n" = 1 ]; then timeout 400 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r3f_c2_${n}gpu.json 2> gpurun_out/r3f_c2_${n}gpu.err
    else timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29530 + n)) bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r3f_c2_${n}gpu.json 2> gpurun_out/r3f_c2_${n}gpu.err; fi
    echo "N=$n rc=$?"; }
timeout 300 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r3f_c2_1gpu_reference.json 2>/dev/null; echo "ref rc=$?"
for n in 1 2 4 8; do run $n; done
python - <<'PY'
import json
base = None
for n in (1, 2, 4, 8):
    try:
        d = json.loads(open(f"gpurun_out/r3f_c2_{n}gpu.json").read().strip().splitlines()[-1])
        base = base or d["value"]
        print("N", n, round(d["value"]/1e6, 2), "M", "eff", round(d["value"]/(n*base), 4), "frac", d["roofline"]["frac"], d["roofline"]["frac_step"], "e2e", round(d["e2e"]["value"]/1e6, 2), "const", round(d["e2e_constants_only"]["value"]/1e6, 2), d["clocks"]["sm_mhz"], d["clocks"]["samples"])
    except Exception as e:
        print("N", n, "unreadable", e)
r = json.loads(open("gpurun_out/r3f_c2_1gpu_reference.json").read().strip().splitlines()[-1]); print("reference", round(r["value"]), r["cpu_baseline"]["cores"])
PY
