mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
tr() { n=$1; shift; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700+n)) bench.py --gpus $n --steps 200 --warmup 20 --no-cpu-baseline "$@" 2>>gpurun_out/r1s.err | tail -1; }
show() { python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(sys.argv[1], d['n_gpus'], 'value %.4e'%d['value'], 'ms/step %.4f'%d['ms_per_step'], 'frac', d['roofline']['frac'], d.get('global_stats'), 'e2e', d.get('e2e',{}).get('value'))" "$1"; }
python bench.py --no-cpu-baseline --no-e2e | show c2-1gpu
tr 2 --no-e2e | tee gpurun_out/r1s_c2_2gpu.json | show c2-bucket8
tr 2 --no-e2e --stats-bucket 1 | show c2-bucket1
tr 2 --no-e2e --dag dag16 --worlds 1024 --entities 1024 | tee gpurun_out/r1s_c5_2gpu.json | show c5-bucket8
tr 2 | tee gpurun_out/r1s_c2_2gpu_e2e.json | show c2-e2e
python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-e2e --worlds 592 --entities 65536 --depth 8 | show 592x64k-auto-team
tail -5 gpurun_out/r1s.err
