#!/bin/bash
# one ticket for both culls of the 16-system set: parity of the DAG cases, then C5 / dag16 4096 bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "dag or ring or tail or mixed" 2>&1 | tail -4
run() { tag=$1; shift; timeout 300 python bench.py --no-e2e --no-cpu-baseline --steps 200 --warmup 20 "$@" 2>>gpurun_out/r1zi.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$tag', 'ms/step %.4f'%d['ms_per_step'], 'frac', d['roofline']['frac'])"; }
run c5 --dag dag16 --worlds 1024
run c5b --dag dag16 --worlds 1024
run dag16-4096 --dag dag16
tail -3 gpurun_out/r1zi.err
