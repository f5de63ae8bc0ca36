#!/bin/bash
# 4 GPUs: default line (staged uploads in e2e), dag8 with the tail grid next to the stats all-reduce stream
mkdir -p gpurun_out
tr() { n=$1; shift; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700+n)) bench.py --gpus $n "$@" 2>>gpurun_out/r1zg.err | tail -1; }
show() { python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(sys.argv[1], d['n_gpus'], 'value %.4e'%d['value'], 'frac', d['roofline']['frac'], d.get('global_stats'), 'e2e', d.get('e2e',{}).get('value'), 'launches', d['gpu_launches'])" "$1"; }
tr 4 --steps 200 --warmup 20 --no-cpu-baseline | tee gpurun_out/r1zg_c2_4gpu.json | show c2
tr 4 --steps 200 --warmup 20 --no-cpu-baseline --no-e2e --dag dag8 | tee gpurun_out/r1zg_dag8_4gpu.json | show dag8
python bench.py --steps 200 --warmup 20 --no-cpu-baseline --no-e2e --dag dag8 | show dag8-1gpu
grep -v "OMP_NUM_THREADS\|\*\*\*\*" gpurun_out/r1zg.err | tail -3
