# two-world interleave in team mode: parity + C3 shapes
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
run() { timeout 300 python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-e2e "$@" 2>>gpurun_out/r1m.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print(' '.join(sys.argv[1:]), '| ms/step %.4f'%d['ms_per_step'], 'kernel_ms', r['kernel_ms'], 'GB/s', r['achieved'], 'frac', r['frac'], '| launches/step', d['gpu_launches']//d['steps'])" "$@"; }
for d in 1 2 4 8; do run --worlds 256 --entities 65536 --depth $d | tee -a gpurun_out/r1m_c3.txt; done
run --dag dag8 --worlds 256 --entities 65536 --depth 8 | tee -a gpurun_out/r1m_c3.txt
for t in 8 16 32; do FIMO_B200_TEAM=$t run --worlds 256 --entities 65536 --depth 8 | sed "s/^/TEAM=$t /" | tee -a gpurun_out/r1m_c3.txt; done
for t in 4 8 16; do FIMO_B200_TEAM=$t run --worlds 592 --entities 65536 --depth 8 | sed "s/^/TEAM=$t /" | tee -a gpurun_out/r1m_c3.txt; done
FIMO_B200_TEAM=8 run --worlds 2048 --entities 8192 --depth 8 | sed "s/^/TEAM=8 /" | tee -a gpurun_out/r1m_c3.txt
run --worlds 2048 --entities 8192 --depth 8 | tee -a gpurun_out/r1m_c3.txt
run --worlds 4096 --entities 1024 --depth 4 | tee -a gpurun_out/r1m_c3.txt
tail -3 gpurun_out/r1m.err
