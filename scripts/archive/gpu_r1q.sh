mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
python bench.py > gpurun_out/bench_r1q.json 2> gpurun_out/bench_r1q.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench_r1q.json')); print(d['value'], d['roofline']['frac']); e=d['e2e']; print({k:e[k] for k in ('value','ms_per_step','sync_value','transfer_value','locked_value')})"; tail -5 gpurun_out/bench_r1q.err
FIMO_B200_IO_MEMCPY=1 python bench.py --no-cpu-baseline | python -c "
import json,sys; d=json.loads(sys.stdin.read()); e=d['e2e']; print('MEMCPY', {k:e[k] for k in ('value','ms_per_step','sync_value','transfer_value','locked_value')})"
