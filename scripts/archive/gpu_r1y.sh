mkdir -p gpurun_out
run() { timeout 300 python bench.py --steps 64 --warmup 16 --no-cpu-baseline --no-e2e "$@" 2>gpurun_out/r1y_last.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print(' '.join(sys.argv[1:]), '| ms/step %.4f'%d['ms_per_step'], 'kernel_ms', r['kernel_ms'], 'frac', r['frac'])" "$@"; grep timing gpurun_out/r1y_last.err | tail -1; }
FIMO_B200_TIMING=1 run --no-graph
FIMO_B200_TIMING=1 run --no-graph --worlds 4096 --entities 4096
FIMO_B200_TIMING=1 run --no-graph --worlds 256 --entities 65536 --depth 1
FIMO_B200_TIMING=1 run --no-graph --worlds 256 --entities 65536 --depth 8
FIMO_B200_TIMING=1 FIMO_B200_PREFETCH=0 run --no-graph --worlds 256 --entities 65536 --depth 8
FIMO_B200_TIMING=1 run --no-graph --dag dag8 --worlds 256 --entities 65536 --depth 8
FIMO_B200_TIMING=1 run --no-graph --dag dag16 --worlds 1024 --entities 1024
FIMO_B200_TIMING=1 run --no-graph --dag dag8
