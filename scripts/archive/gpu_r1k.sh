mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
python bench.py > gpurun_out/bench_r1k.json 2> gpurun_out/bench_r1k.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench_r1k.json')); print(d['value'], d['roofline']['frac']); print(json.dumps(d['e2e'], indent=1))"; tail -5 gpurun_out/bench_r1k.err
