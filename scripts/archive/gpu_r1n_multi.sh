# 8-GPU box: BASELINE configs[3] (32768 worlds x 4k entities sharded over 8 GPUs) and configs[4] (dag16, 1M entities/GPU) at 2/4/8
mkdir -p gpurun_out
tr() { n=$1; shift; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) bench.py --gpus $n --steps 100 --warmup 10 --no-cpu-baseline "$@" 2>>gpurun_out/r1n.err | tail -1; }
tr 8 --worlds 4096 --entities 4096 --depth 4 --no-e2e > gpurun_out/r1n_c4_8gpu.json; cat gpurun_out/r1n_c4_8gpu.json | cut -c1-600
for n in 2 4 8; do tr $n --dag dag16 --worlds 1024 --entities 1024 --depth 4 --no-e2e > gpurun_out/r1n_c5_${n}gpu.json; cut -c1-200 gpurun_out/r1n_c5_${n}gpu.json; done
timeout 200 python bench.py --dag dag16 --worlds 1024 --entities 1024 --depth 4 --steps 100 --warmup 10 --no-cpu-baseline --no-e2e 2>>gpurun_out/r1n.err | tail -1 > gpurun_out/r1n_c5_1gpu.json; cut -c1-200 gpurun_out/r1n_c5_1gpu.json
tr 8 > gpurun_out/r1n_c2_8gpu.json; cut -c1-200 gpurun_out/r1n_c2_8gpu.json
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r1n_*.json')):
    try:
        d=json.load(open(f)); print(f, d['n_gpus'], 'value %.3e'%d['value'], 'ms/step %.4f'%d['ms_per_step'], 'frac', d['roofline']['frac'], d.get('global_stats'))
    except Exception as e: print(f, 'ERR', e)
PY
tail -5 gpurun_out/r1n.err
