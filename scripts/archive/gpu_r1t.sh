mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
run() { timeout 300 python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-e2e "$@" 2>>gpurun_out/r1t.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print(' '.join(sys.argv[1:]), '| ms/step %.4f'%d['ms_per_step'], 'kernel_ms', r['kernel_ms'], 'GB/s', r['achieved'], 'frac', r['frac'], '| launches/step', d['gpu_launches']//d['steps'])" "$@"; }
for p in "" 128,4,4 128,3,4 128,3,5 128,3,6 256,2,3 256,2,4 128,2,8; do
  FIMO_B200_PIPE=$p run --worlds 256 --entities 65536 --depth 8 | sed "s/^/PIPE=$p /" | tee -a gpurun_out/r1t_c3.txt
done
for p in "" 128,4,4 128,3,4 128,3,5; do
  FIMO_B200_PIPE=$p run --dag dag8 --worlds 256 --entities 65536 --depth 8 | sed "s/^/PIPE=$p /" | tee -a gpurun_out/r1t_c3.txt
done
tail -3 gpurun_out/r1t.err
