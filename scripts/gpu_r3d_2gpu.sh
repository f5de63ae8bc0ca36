python -m pytest tests/test_multi_gpu.py -m gpu -q > gpurun_out/r3d_pytest_2gpu.log 2>&1; echo mgpu rc=$?; tail -2 gpurun_out/r3d_pytest_2gpu.log
for S in peers nccl; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 20 --warmup 5 --stats $S > gpurun_out/r3d_bench_c2_2gpu_$S.json 2> gpurun_out/r3d_bench_2gpu_$S.err; echo bench $S rc=$?
  tail -c 600 gpurun_out/r3d_bench_c2_2gpu_$S.json | head -c 0
  python - <<PY
import json
d=json.loads(open("gpurun_out/r3d_bench_c2_2gpu_$S.json").read().strip().splitlines()[-1])
print(round(d["value"]/1e6,2), round(d["e2e"]["value"]/1e6,2), d["roofline"]["frac"], str(d.get("global_stats"))[:300])
PY
done
