# C3 shape sensitivity: hierarchy depth (hand-offs + gathers), team size, on the 3-system chain and dag8
mkdir -p gpurun_out
run() { python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-e2e "$@" 2>>gpurun_out/r1l.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print(' '.join(sys.argv[1:]), '| ms/step %.4f'%d['ms_per_step'], 'kernel_ms', r['kernel_ms'], 'GB/s', r['achieved'], 'frac', r['frac'], '| launches/step', d['gpu_launches']//d['steps'])" "$@"; }
for d in 1 2 4 8; do run --worlds 256 --entities 65536 --depth $d | tee -a gpurun_out/r1l_c3.txt; done
for t in 4 8 32; do FIMO_B200_TEAM=$t run --worlds 256 --entities 65536 --depth 8 | sed "s/^/TEAM=$t /" | tee -a gpurun_out/r1l_c3.txt; done
FIMO_B200_PREFETCH=0 run --worlds 256 --entities 65536 --depth 8 | sed "s/^/PREFETCH=0 /" | tee -a gpurun_out/r1l_c3.txt
FIMO_B200_PUBLISH=chunk run --worlds 256 --entities 65536 --depth 8 | sed "s/^/PUBLISH=chunk /" | tee -a gpurun_out/r1l_c3.txt
run --worlds 592 --entities 65536 --depth 8 | tee -a gpurun_out/r1l_c3.txt
FIMO_B200_TEAM=1 run --worlds 592 --entities 65536 --depth 8 | sed "s/^/TEAM=1 /" | tee -a gpurun_out/r1l_c3.txt
tail -3 gpurun_out/r1l.err
