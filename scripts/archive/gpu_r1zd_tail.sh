#!/bin/bash
# tail grid experiments: worlds per tail CTA, single-round batches
mkdir -p gpurun_out
run() { tag=$1; shift; timeout 300 python bench.py --no-e2e --no-cpu-baseline --steps 200 --warmup 20 "$@" 2>>gpurun_out/r1zd.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$tag', 'ms/step %.4f'%d['ms_per_step'], 'frac', d['roofline']['frac'], 'launches', d['gpu_launches'])"; }
FIMO_B200_TAIL_WORLDS=2 run c2-tw2
FIMO_B200_TAIL_WORLDS=2 run dag8-tw2 --dag dag8
FIMO_B200_TAIL_WORLDS=2 FIMO_B200_TAIL=2 run dag8-tw2-t2 --dag dag8
FIMO_B200_TAIL_WORLDS=2 run dag16-tw2 --dag dag16
FIMO_B200_TAIL=0 run c5-off --dag dag16 --worlds 1024
FIMO_B200_TAIL_MIN_ROUNDS=1 run c5-r1 --dag dag16 --worlds 1024
FIMO_B200_TAIL=0 run dag8-1024-off --dag dag8 --worlds 1024
FIMO_B200_TAIL_MIN_ROUNDS=1 run dag8-1024-r1 --dag dag8 --worlds 1024
FIMO_B200_TAIL=0 run dag8-2048-off --dag dag8 --worlds 2048
run dag8-2048-on --dag dag8 --worlds 2048
FIMO_B200_TAIL=0 run dag8-8192-off --dag dag8 --worlds 8192
run dag8-8192-on --dag dag8 --worlds 8192
tail -3 gpurun_out/r1zd.err
