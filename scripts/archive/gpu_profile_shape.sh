# ncu launch list + one full capture of the world-step kernel for a given shape: SHAPE="--worlds .. --entities .. --depth .." NAME=tag
mkdir -p gpurun_out
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e $SHAPE > gpurun_out/plain_$NAME.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_$NAME.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e $SHAPE > gpurun_out/ncu1_$NAME.log 2>&1
echo "ncu list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_world_step_pipe -s 3 -c 1 -o gpurun_out/prof_$NAME python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e $SHAPE > gpurun_out/ncu2_$NAME.log 2>&1
echo "ncu full rc=$?"
