# 8-GPU box, statistics ring + bucketed all-reduce: configs[3], configs[4] sweep, configs[1] shape per GPU
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
tr() { n=$1; shift; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29800+n)) bench.py --gpus $n --steps 200 --warmup 20 --no-cpu-baseline "$@" 2>>gpurun_out/r1x.err | tail -1; }
tr 8 --worlds 4096 --entities 4096 --depth 4 --no-e2e > gpurun_out/r1x_c4_8gpu.json
for n in 2 4 8; do tr $n --dag dag16 --worlds 1024 --entities 1024 --depth 4 --no-e2e > gpurun_out/r1x_c5_${n}gpu.json; done
timeout 200 python bench.py --dag dag16 --worlds 1024 --entities 1024 --depth 4 --no-cpu-baseline --no-e2e 2>>gpurun_out/r1x.err | tail -1 > gpurun_out/r1x_c5_1gpu.json
for n in 2 4 8; do tr $n --no-e2e > gpurun_out/r1x_c2_${n}gpu.json; done
tr 8 > gpurun_out/r1x_c2_8gpu_e2e.json
timeout 200 python bench.py --no-cpu-baseline --no-e2e 2>>gpurun_out/r1x.err | tail -1 > gpurun_out/r1x_c2_1gpu.json
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r1x_*.json')):
    try:
        d=json.load(open(f)); print(f, d['n_gpus'], 'value %.4e'%d['value'], 'ms/step %.4f'%d['ms_per_step'], 'frac', d['roofline']['frac'], d.get('global_stats'), 'e2e', d.get('e2e',{}).get('value'))
    except Exception as e: print(f, 'ERR', e)
PY
grep -v "OMP_NUM\|^\*" gpurun_out/r1x.err | tail -5
