#!/bin/bash
# Geometry sweep of the one-CTA-per-world kernel with generations sharing a launch (C2 and the C4 shape).
mkdir -p gpurun_out
S="--steps 32 --warmup 16 --windows 3 --no-cpu-baseline --no-e2e --no-other-configs --no-parity-gate"
run() { tag=$1; pipe=$2; shift 2; FIMO_B200_PIPE=$pipe timeout 100 python bench.py $S "$@" 2>/dev/null | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.read().strip().splitlines()[-1]); r = d['roofline']
    print('$tag', 'PIPE=$pipe', round(d['value'] / 1e6, 3), 'M', 'frac', r['frac'], 'launches', d['gpu_launches'], r['kernel'][:100])
except Exception as e:
    print('$tag', 'PIPE=$pipe', 'failed', e)
"; }
{
for p in "" 128,4,3 128,5,3 128,5,2 128,6,2 128,4,4 256,2,3 256,3,3 256,3,2; do run c2 "$p"; done
for p in "" 256,2,3 256,3,3 256,3,2 128,4,3 512,1,3; do run c4 "$p" --entities 4096; done
for p in "" 128,4,3 128,4,2 256,2,3; do run dag16 "$p" --dag dag16 --worlds 1024; done
} > gpurun_out/r2v_sweep.txt 2>&1
cat gpurun_out/r2v_sweep.txt
