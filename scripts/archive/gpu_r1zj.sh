#!/bin/bash
# e2e with the generation as a CUDA graph launch vs plain stream launches between the copy kernels
mkdir -p gpurun_out
run() { tag=$1; shift; timeout 300 python bench.py --no-cpu-baseline --steps 200 --warmup 20 "$@" 2>>gpurun_out/r1zj.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); e=d['e2e']; print('$tag', 'ms/step %.4f'%d['ms_per_step'], 'frac', d['roofline']['frac'], 'e2e %.4e'%e['value'], 'e2e ms %.4f'%e['ms_per_step'], 'sync %.4e'%e['sync_value'])"; }
run graph
run nograph --no-graph
run graph2
run nograph2 --no-graph
run dag16-graph --dag dag16 --worlds 1024
run dag16-nograph --dag dag16 --worlds 1024 --no-graph
tail -3 gpurun_out/r1zj.err
