# ncu evidence for the current kernels: launch lists of the bench command (C2) and of dag16 at 1M entities, one
# --set full capture of the world-step kernel for each.  Every ncu command runs only after the same command exited 0 plain.
mkdir -p gpurun_out
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e"
$B > gpurun_out/p_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_r1p_c2.csv $B > gpurun_out/ncu1.log 2>&1; echo "c2 list rc=$?"
$B > gpurun_out/p_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_world_step_pipe -s 3 -c 1 -o gpurun_out/prof_r1p_c2 $B > gpurun_out/ncu2.log 2>&1; echo "c2 full rc=$?"
D="$B --dag dag16 --worlds 1024 --entities 1024 --depth 4"
$D > gpurun_out/p_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_r1p_dag16.csv $D > gpurun_out/ncu3.log 2>&1; echo "dag16 list rc=$?"
$D > gpurun_out/p_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_world_step_pipe -s 3 -c 1 -o gpurun_out/prof_r1p_dag16 $D > gpurun_out/ncu4.log 2>&1; echo "dag16 full rc=$?"
E="$B --dag dag8 --worlds 256 --entities 65536 --depth 8"
$E > gpurun_out/p_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_world_step_pipe -s 3 -c 1 -o gpurun_out/prof_r1p_c3dag8 $E > gpurun_out/ncu5.log 2>&1; echo "c3 full rc=$?"
ls -la gpurun_out | tail -12
