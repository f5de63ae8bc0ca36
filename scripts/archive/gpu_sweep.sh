# Sweeps the fused world-step launch plans (FIMO_B200_PIPE=T,blocks,stages[,cap]) on the B200 box.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for cfg in ${SWEEP:-128,4,4 256,2,4}; do
  FIMO_B200_PIPE=$cfg python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-e2e ${BENCH_ARGS} 2> gpurun_out/sweep_$cfg.err \
    | python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print('$cfg', 'ms/step %.5f'%d['ms_per_step'], 'kernel_ms', r['kernel_ms'], 'GB/s', r['achieved'], 'frac', r['frac'])"
done
for cfg in ${PARITY:-128,4,4 512,1,4 256,2,4,100}; do
  echo "parity under FIMO_B200_PIPE=$cfg"; FIMO_B200_PIPE=$cfg timeout 600 python -m pytest tests -m gpu -x -q -k "three_system or large_world or full_size" 2>&1 | tail -2
done
