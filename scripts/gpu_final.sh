# final verification of main: GPU tests, smoke, default bench (refreshes profiles), reference arm
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench_final.json')); print(d['value'], d['roofline']['frac'], d['e2e']['value'], d['cpu_baseline']['value'], d['clocks'], d['gpu_launches'])"
python bench.py --dag dag8 --worlds 256 --entities 65536 --depth 8 --no-cpu-baseline --no-e2e > gpurun_out/bench_final_c3.json 2>>gpurun_out/bench_final.err
python bench.py --dag dag16 --worlds 1024 --entities 1024 --no-cpu-baseline --no-e2e > gpurun_out/bench_final_c5.json 2>>gpurun_out/bench_final.err
python -c "
import json
for f in ('c3','c5'):
    d=json.load(open('gpurun_out/bench_final_%s.json'%f)); print(f, d['value'], d['ms_per_step'], d['roofline']['frac'], d['gpu_launches']//d['steps'])"
