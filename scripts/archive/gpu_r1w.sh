mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/bench_r1w.json 2> gpurun_out/bench_r1w.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench_r1w.json')); print(d['value'], d['roofline']['frac'], d['e2e']['value'], d['cpu_baseline']['value'], d['clocks'])"
python bench.py --impl reference --steps 20 --warmup 3 | cut -c1-400
