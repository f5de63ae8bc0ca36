#!/bin/bash
# pacing experiment: early-placed CTAs of an SM wait before refilling a stage (FIMO_B200_PACE = ns for slot 0)
mkdir -p gpurun_out
run() { tag=$1; shift; timeout 300 python bench.py --no-e2e --no-cpu-baseline --steps 200 --warmup 20 "$@" 2>>gpurun_out/r1zf.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$tag', 'ms/step %.4f'%d['ms_per_step'], 'frac', d['roofline']['frac'])"; }
for p in 0 60 120 200 300 450; do FIMO_B200_PACE=$p run c2-pace$p; done
for p in 0 150 300; do FIMO_B200_PACE=$p run dag16-pace$p --dag dag16; done
for p in 0 150 300; do FIMO_B200_PACE=$p FIMO_B200_TAIL=0 run dag8-notail-pace$p --dag dag8; done
for p in 150 300; do FIMO_B200_PACE=$p run dag8-tail-pace$p --dag dag8; done
for p in 0 150 300; do FIMO_B200_PACE=$p run c5-pace$p --dag dag16 --worlds 1024; done
tail -3 gpurun_out/r1zf.err
