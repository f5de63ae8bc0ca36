mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python bench.py > gpurun_out/bench_r1i.json 2> gpurun_out/bench_r1i.err; echo "bench rc=$?"; tail -c 3000 gpurun_out/bench_r1i.json; tail -5 gpurun_out/bench_r1i.err
python bench.py --no-fusion --no-cpu-baseline --no-e2e > gpurun_out/bench_r1i_unfused.json 2>&1; tail -c 1500 gpurun_out/bench_r1i_unfused.json
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_r1i.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu1.log 2>&1
echo "ncu list rc=$?"
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_world_step_pipe -s 3 -c 1 -o gpurun_out/prof_r1i python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu2.log 2>&1
echo "ncu full rc=$?"; ls -la gpurun_out
