mkdir -p gpurun_out
run() { timeout 300 python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-e2e "$@" 2>>gpurun_out/r1u.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print(' '.join(sys.argv[1:]), '| ms/step %.4f'%d['ms_per_step'], 'kernel_ms', r['kernel_ms'], 'GB/s', r['achieved'], 'frac', r['frac'], '| launches/step', d['gpu_launches']//d['steps'])" "$@"; }
for cfg in "128,3,4 16" "128,3,4 12" "128,2,6 16" "128,2,8 8" "128,2,8 16" "256,2,3 16" "256,2,4 8" "256,2,3 8" "256,1,6 8" "512,1,3 8" "128,4,3 12" "128,4,3 20"; do
  set -- $cfg
  FIMO_B200_PIPE=$1 FIMO_B200_TEAM=$2 run --worlds 256 --entities 65536 --depth 8 | sed "s/^/PIPE=$1 TEAM=$2 /" | tee -a gpurun_out/r1u_c3.txt
done
tail -3 gpurun_out/r1u.err
