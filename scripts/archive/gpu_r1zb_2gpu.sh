#!/bin/bash
# 2 GPUs with staged uploads: default line (with e2e) and the reference arm under torchrun
mkdir -p gpurun_out
tr() { n=$1; shift; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700+n)) bench.py --gpus $n "$@" 2>>gpurun_out/r1zb.err | tail -1; }
tr 2 --steps 200 --warmup 20 --no-cpu-baseline | tee gpurun_out/r1zb_c2_2gpu.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['n_gpus'], 'value %.4e'%d['value'], 'frac', d['roofline']['frac'], d.get('global_stats'), 'e2e', d['e2e']['value'], d['e2e'].get('sync_value'))"
tr 2 --impl reference --steps 2 --warmup 1 | tee gpurun_out/r1zb_ref_2gpu.json | cut -c1-300
tail -3 gpurun_out/r1zb.err
