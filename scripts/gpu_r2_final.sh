#!/bin/bash
# Round-2 final verification on one B200: the GPU suite, the driver's bench command and reference arm, the other BASELINE shapes,
# and the ncu evidence (launch lists of the bench command + --set full captures of the dominant kernel) for the final tree.
# Every command has its own timeout; every ncu command runs only after the same command exited 0 without ncu.
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -x -q --timeout 100 --timeout-method=thread > gpurun_out/r2w_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2w_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2w_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r2w_smoke.log
timeout 400 python bench.py --steps 20 --warmup 5 > gpurun_out/r2w_bench_c2.json 2> gpurun_out/r2w_bench_c2.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2w_bench_c2_reference_arm.json 2>/dev/null; echo "ref rc=$?"
S="--steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-other-configs"
timeout 120 python bench.py $S --entities 4096 > gpurun_out/r2w_bench_c4_shape.json 2>/dev/null; echo "c4 rc=$?"
timeout 120 python bench.py $S --dag dag16 --worlds 1024 > gpurun_out/r2w_bench_c5_shape.json 2>/dev/null; echo "c5 rc=$?"
timeout 120 python bench.py $S --dag dag8 --worlds 256 --entities 65536 --depth 8 > gpurun_out/r2w_bench_c3.json 2>/dev/null; echo "c3 rc=$?"
timeout 120 python bench.py $S --dag dag8 > gpurun_out/r2w_bench_dag8_4096.json 2>/dev/null; echo "dag8 rc=$?"
timeout 120 python bench.py $S --dag dag16 > gpurun_out/r2w_bench_dag16_4096.json 2>/dev/null; echo "dag16 rc=$?"
for f in gpurun_out/r2w_bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r = d.get("roofline", {})
    print(sys.argv[1].split("/")[-1], round(d["value"] / 1e6, 3), "M", "frac", r.get("frac"), "frac_step", r.get("frac_step"), "e2e", (d.get("e2e") or {}).get("value"),
          "launches", d.get("gpu_launches"), "one-gen", (r.get("one_generation_per_launch") or {}).get("frac"), "cpu", (d.get("cpu_baseline") or {}).get("value"), d.get("clocks"))
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
export FIMO_B200_GENS_HOLD=16
B="timeout 200 python bench.py --steps 32 --warmup 16 --windows 1 --no-cpu-baseline --no-e2e --no-other-configs --no-parity-gate"
run() { # tag, extra bench args...
    tag=$1; shift
    $B "$@" > gpurun_out/plain_r2w_$tag.json 2> gpurun_out/plain_r2w_$tag.err || { echo "$tag plain failed"; tail -3 gpurun_out/plain_r2w_$tag.err; return; }
    timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_r2w_$tag.csv $B "$@" > gpurun_out/ncu_list_r2w_$tag.log 2>&1; echo "$tag list rc=$?"
    timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_world_step_pipe -c 2 -f -o gpurun_out/prof_r2w_$tag $B "$@" > gpurun_out/ncu_full_r2w_$tag.log 2>&1; echo "$tag full rc=$?"
}
run c2
run c5 --dag dag16 --worlds 1024
run c3 --dag dag8 --worlds 256 --entities 65536 --depth 8
