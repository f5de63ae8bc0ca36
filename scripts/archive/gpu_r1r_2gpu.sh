mkdir -p gpurun_out
tr() { n=$1; shift; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n)) bench.py --gpus $n --steps 200 --warmup 20 --no-cpu-baseline "$@" 2>>gpurun_out/r1r.err | tail -1; }
show() { python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(sys.argv[1], d['n_gpus'], 'value %.4e'%d['value'], 'ms/step %.4f'%d['ms_per_step'], 'frac', d['roofline']['frac'], d.get('global_stats'), 'e2e', d.get('e2e',{}).get('value'))" "$1"; }
tr 2 --no-e2e | tee gpurun_out/r1r_c2_2gpu.json | show c2-bucket8
tr 2 --no-e2e --stats-bucket 1 | show c2-bucket1
tr 2 --no-e2e --dag dag16 --worlds 1024 --entities 1024 | tee gpurun_out/r1r_c5_2gpu.json | show c5-bucket8
tr 2 --no-e2e --dag dag16 --worlds 1024 --entities 1024 --stats-bucket 1 | show c5-bucket1
tr 2 | tee gpurun_out/r1r_c2_2gpu_e2e.json | show c2-e2e
tail -5 gpurun_out/r1r.err
