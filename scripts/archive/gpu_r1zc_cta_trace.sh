#!/bin/bash
# TIMING build only: per-CTA (smid, entry, loop start, loop end, exit) trace of the one-CTA-per-world kernel
mkdir -p gpurun_out
run() { tag=$1; shift; FIMO_B200_TIMING=1 FIMO_B200_TIMING_DUMP=gpurun_out/r1ze_trace_$tag.csv timeout 300 python bench.py --no-graph --no-e2e --no-cpu-baseline --steps 64 --warmup 16 "$@" 2>gpurun_out/r1ze_$tag.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$tag', 'ms/step %.4f'%d['ms_per_step'], 'frac', d['roofline']['frac'])"; grep timing gpurun_out/r1ze_$tag.err | tail -1; }
run dag16_4096 --dag dag16
FIMO_B200_TAIL=1 run dag16_4096_tail --dag dag16
FIMO_B200_TAIL=0 run dag8_notail --dag dag8
