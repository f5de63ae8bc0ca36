#!/bin/bash
# Round-2, one 8-GPU box: the driver's scaling sweep (N = 1, 2, 4, 8 back to back, --steps 20 --warmup 5) for configs[1]'s shape,
# then configs[3] (32768 worlds x 4k entities over 8 GPUs) and configs[4] (dag16, 1M entities per GPU) at 8.
mkdir -p gpurun_out
run() { # N, tag, extra args
    n=$1; tag=$2; shift 2
    if [ "$n" = 1 ]; then python bench.py --gpus 1 --steps 20 --warmup 5 "$@" > gpurun_out/r2k_${tag}_${n}gpu.json 2> gpurun_out/r2k_${tag}_${n}gpu.err
    else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29520 + n)) bench.py --gpus $n --steps 20 --warmup 5 "$@" > gpurun_out/r2k_${tag}_${n}gpu.json 2> gpurun_out/r2k_${tag}_${n}gpu.err; fi
    echo "$tag N=$n rc=$?"
}
for n in 1 2 4 8; do run $n c2 --no-cpu-baseline --no-other-configs; done
run 8 c4 --entities 4096 --no-e2e
run 8 c5 --dag dag16 --worlds 1024 --no-e2e
run 8 c2nccl --stats nccl --no-e2e
for f in gpurun_out/r2k_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r = d.get("roofline", {})
    print(sys.argv[1].split("/")[-1], "N", d["n_gpus"], round(d["value"] / 1e6, 3), "M", "frac", r.get("frac"), "frac_step", r.get("frac_step"), "e2e", (d.get("e2e") or {}).get("value"),
          "launches", d.get("gpu_launches"), "clocks", d.get("clocks", {}).get("sm_mhz"))
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
