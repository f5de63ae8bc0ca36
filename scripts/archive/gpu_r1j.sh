# round-1 session j: parity (all GPU tests), bench (C2 with pipelined e2e), dag16 fused set 2 vs set 1, dag8 C3
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
python bench.py > gpurun_out/bench_r1j.json 2> gpurun_out/bench_r1j.err; echo "bench rc=$?"; tail -c 3500 gpurun_out/bench_r1j.json; tail -5 gpurun_out/bench_r1j.err
run() { python bench.py --steps 100 --warmup 5 --no-cpu-baseline --no-e2e "$@" 2>>gpurun_out/r1j.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print(' '.join(sys.argv[1:]), '| ms/step %.4f'%d['ms_per_step'], 'kernel_ms', r['kernel_ms'], 'GB/s', r['achieved'], 'frac', r['frac'], '| launches/step', d['gpu_launches']//d['steps'])" "$@"; }
run --dag dag16 --worlds 1024 --entities 1024 --depth 4 | tee -a gpurun_out/r1j_dags.txt
FIMO_B200_NO_SET2=1 run --dag dag16 --worlds 1024 --entities 1024 --depth 4 | sed 's/^/NO_SET2 /' | tee -a gpurun_out/r1j_dags.txt
run --dag dag16 --worlds 4096 --entities 1024 --depth 4 | tee -a gpurun_out/r1j_dags.txt
FIMO_B200_NO_SET2=1 run --dag dag16 --worlds 4096 --entities 1024 --depth 4 | sed 's/^/NO_SET2 /' | tee -a gpurun_out/r1j_dags.txt
FIMO_B200_PIPE=128,3,3 run --dag dag16 --worlds 4096 --entities 1024 --depth 4 | sed 's/^/PIPE=128,3,3 /' | tee -a gpurun_out/r1j_dags.txt
FIMO_B200_PIPE=256,2,3 run --dag dag16 --worlds 4096 --entities 1024 --depth 4 | sed 's/^/PIPE=256,2,3 /' | tee -a gpurun_out/r1j_dags.txt
run --dag dag8 --worlds 256 --entities 65536 --depth 8 | tee -a gpurun_out/r1j_dags.txt
run --dag dag8 --worlds 4096 --entities 1024 --depth 4 | tee -a gpurun_out/r1j_dags.txt
tail -5 gpurun_out/r1j.err
