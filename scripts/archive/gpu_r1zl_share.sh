#!/bin/bash
# weighted static share by placement order (FIMO_B200_SHARE = relative time per chunk of the 1st..4th CTA of an SM, in 1/1000)
mkdir -p gpurun_out
FIMO_B200_SHARE=919,967,1023,1091 timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "full_size" 2>&1 | tail -3
run() { tag=$1; shift; timeout 300 python bench.py --no-e2e --no-cpu-baseline --steps 200 --warmup 20 "$@" 2>>gpurun_out/r1zl.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$tag', 'ms/step %.4f'%d['ms_per_step'], 'frac', d['roofline']['frac'])"; }
run dag16-off --dag dag16
FIMO_B200_SHARE=919,967,1023,1091 run dag16-share --dag dag16
FIMO_B200_SHARE=880,950,1030,1140 run dag16-share-strong --dag dag16
run c5-off --dag dag16 --worlds 1024
FIMO_B200_SHARE=919,967,1023,1091 run c5-share --dag dag16 --worlds 1024
run dag16-8192-off --dag dag16 --worlds 8192
FIMO_B200_SHARE=919,967,1023,1091 run dag16-8192-share --dag dag16 --worlds 8192
run c2-off
FIMO_B200_SHARE=950,983,1014,1053 run c2-share
run c2-8192-off --worlds 8192
FIMO_B200_SHARE=950,983,1014,1053 run c2-8192-share --worlds 8192
tail -3 gpurun_out/r1zl.err
