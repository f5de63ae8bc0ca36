# Roofline of the fused world-step on the BASELINE world shapes (1 GPU): C2, C3-shape, C4-shape, small batch.
mkdir -p gpurun_out
for a in "--worlds 4096 --entities 1024 --depth 4" "--worlds 256 --entities 65536 --depth 8" "--worlds 4096 --entities 4096 --depth 4" "--worlds 16 --entities 65536 --depth 8" "--worlds 2048 --entities 8192 --depth 8" "--worlds 1024 --entities 1024 --depth 4" ${EXTRA_SHAPES}; do
  python bench.py --steps ${STEPS:-50} --warmup 5 --no-cpu-baseline --no-e2e $a 2>>gpurun_out/shapes.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print('$a', '| ms/step %.4f'%d['ms_per_step'], 'kernel_ms', r['kernel_ms'], 'GB/s', r['achieved'], 'frac', r['frac'], '| launches/step', d['gpu_launches']//d['steps'])"
done
tail -5 gpurun_out/shapes.err
