#!/bin/bash
# Plan sweep under multi-generation launches: does the window-coverage heuristic still pick the best geometry?
mkdir -p gpurun_out
B="python bench.py --steps 64 --warmup 8 --windows 3 --no-cpu-baseline --no-e2e --no-other-configs --no-parity-gate"
one() { # label, env assignments..., -- bench args
    label=$1; shift
    envs=()
    while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
    env "${envs[@]}" $B "$@" 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$label', round(d['value']/1e6,3), 'M', round(d['ms_per_step'],5), 'ms frac', d['roofline']['frac'], 'launches', d['gpu_launches'], '|', d['roofline']['kernel'][:90])"
}
one c4-default X=1 -- --entities 4096
for p in 128,4,3 256,2,3 256,3,3 512,1,3 256,2,2; do one c4-$p FIMO_B200_PIPE=$p -- --entities 4096; done
one c2-default X=1 --
for p in 128,4,2 128,4,4 128,5,3 256,2,3 256,3,3; do one c2-$p FIMO_B200_PIPE=$p --; done
one c2-block2 FIMO_B200_BLOCK=2 --
one c2-gens4 FIMO_B200_GENS=4 --
one c2-gens8 FIMO_B200_GENS=8 --
one dag16-4096-default X=1 -- --dag dag16
for p in 128,4,2 256,2,3 256,3,3; do one dag16-4096-$p FIMO_B200_PIPE=$p -- --dag dag16; done
one c5-default X=1 -- --dag dag16 --worlds 1024
for p in 128,4,2 256,2,3 256,3,3; do one c5-$p FIMO_B200_PIPE=$p -- --dag dag16 --worlds 1024; done
one dag8-default X=1 -- --dag dag8
one dag8-notail FIMO_B200_TAIL=0 -- --dag dag8
