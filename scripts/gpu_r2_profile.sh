#!/bin/bash
# Round-2 ncu evidence: launch lists of the bench command (driver flags) and --set full captures of the dominant kernel for
# C2 (chain3), the configs[4] per-GPU shape (dag16) and configs[2] (dag8, CTA teams).
# Under ncu every launch is synchronous, so generations never queue up behind one another and the product would only ever
# issue single-generation launches; FIMO_B200_GENS_HOLD=16 makes the group hold plain generations until 16 are pending, which
# is the launch shape the un-profiled bench reaches by itself (see gpu_launches in its JSON line).
# Every ncu command runs only after the same command exited 0 without ncu.
mkdir -p gpurun_out
export FIMO_B200_GENS_HOLD=16
B="python bench.py --steps 32 --warmup 16 --windows 1 --no-cpu-baseline --no-e2e --no-other-configs --no-parity-gate"
run() { # tag, extra bench args...
    tag=$1; shift
    $B "$@" > gpurun_out/plain_$tag.json 2> gpurun_out/plain_$tag.err || { echo "$tag plain failed"; tail -3 gpurun_out/plain_$tag.err; return; }
    ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_r2g_$tag.csv $B "$@" > gpurun_out/ncu_list_$tag.log 2>&1; echo "$tag list rc=$?"
    ncu --set full --clock-control none --import-source on -k regex:k_world_step_pipe -c 3 -f -o gpurun_out/prof_r2g_$tag $B "$@" > gpurun_out/ncu_full_$tag.log 2>&1; echo "$tag full rc=$?"
}
run c2
run c5 --dag dag16 --worlds 1024
run c3 --dag dag8 --worlds 256 --entities 65536 --depth 8
ls -la gpurun_out | grep -c prof_r2g
