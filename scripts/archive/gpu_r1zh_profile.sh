#!/bin/bash
# ncu evidence after staged uploads + tail grid: launch list of the bench command with its e2e leg, launch list and one
# --set full capture of the 8-system set at 4096 x 1024 (persistent + tail grid).  Each ncu command only after the same command exited 0 plain.
mkdir -p gpurun_out
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/p_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r1zh_c2_e2e.csv $B > gpurun_out/ncu1.log 2>&1; echo "c2 list rc=$?"
D="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --dag dag8"
$D > gpurun_out/p_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_r1zh_dag8.csv $D > gpurun_out/ncu2.log 2>&1; echo "dag8 list rc=$?"
$D > gpurun_out/p_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_world_step_pipe -s 6 -c 2 -o gpurun_out/prof_r1zh_dag8 $D > gpurun_out/ncu3.log 2>&1; echo "dag8 full rc=$?"
ls -la gpurun_out | grep r1zh
