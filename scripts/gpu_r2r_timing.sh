#!/bin/bash
# TIMING build on the box (make TIMING=1): where the chunk loop's time goes now, team mode (C3 shapes) and plain dag8.
mkdir -p gpurun_out
timeout 400 make -C fimo_b200/csrc -B -s -j16 TIMING=1 > gpurun_out/r2r_make.log 2>&1; echo "make rc=$?"
run() { FIMO_B200_TIMING=1 timeout 200 python bench.py --steps 64 --warmup 16 --windows 1 --no-cpu-baseline --no-e2e --no-other-configs --no-parity-gate --no-graph "$@" 2>gpurun_out/r2r_last.err | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; print(' '.join(sys.argv[1:]), '| ms/step %.4f'%d['ms_per_step'], 'kernel_ms', r['kernel_ms'], 'frac', r['frac'])" "$@"; grep timing gpurun_out/r2r_last.err | tail -1; }
{
run --dag dag8 --worlds 256 --entities 65536 --depth 8
run --worlds 256 --entities 65536 --depth 8
run --worlds 256 --entities 65536 --depth 1
run --dag dag8 --worlds 256 --entities 65536 --depth 2
run --dag dag8 --worlds 256 --entities 65536 --depth 4
FIMO_B200_GENS=1 run --dag dag8
} > gpurun_out/r2r_timing.txt 2>&1
cat gpurun_out/r2r_timing.txt
