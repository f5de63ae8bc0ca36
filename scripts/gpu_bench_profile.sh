mkdir -p gpurun_out
python bench.py > gpurun_out/bench_r1a.json 2> gpurun_out/bench_r1a.err; echo "bench rc=$?"; tail -c 3000 gpurun_out/bench_r1a.json; tail -5 gpurun_out/bench_r1a.err
python bench.py --no-fusion --no-cpu-baseline --no-e2e > gpurun_out/bench_r1a_unfused.json 2>&1; tail -c 1500 gpurun_out/bench_r1a_unfused.json
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_r1a.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu1.log 2>&1
echo "ncu list rc=$?"
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_world_cta -s 3 -c 2 -o gpurun_out/prof_r1a python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu2.log 2>&1
echo "ncu full rc=$?"; ls -la gpurun_out
