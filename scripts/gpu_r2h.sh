#!/bin/bash
# Round-2 call 8: team-local L2-resident rows + L2 eviction priorities (C3), prev_copy loads ahead of the compaction (C5).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2h_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2h_pytest.log
S="--steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-other-configs"
C3="--dag dag8 --worlds 256 --entities 65536 --depth 8"
python bench.py $S $C3 > gpurun_out/r2h_bench_c3.json 2>/dev/null; echo "c3 rc=$?"
FIMO_B200_L2_HINTS=0 python bench.py $S $C3 > gpurun_out/r2h_bench_c3_nohints.json 2>/dev/null; echo "c3 nohints rc=$?"
python bench.py $S --worlds 256 --entities 65536 --depth 8 > gpurun_out/r2h_bench_c3_chain3.json 2>/dev/null; echo "c3 chain3 rc=$?"
python bench.py $S --dag dag8 --worlds 592 --entities 65536 --depth 8 > gpurun_out/r2h_bench_592x64k.json 2>/dev/null; echo "592 rc=$?"
python bench.py $S --dag dag16 --worlds 1024 > gpurun_out/r2h_bench_c5_shape.json 2>/dev/null; echo "c5 rc=$?"
python bench.py $S --dag dag16 > gpurun_out/r2h_bench_dag16_4096.json 2>/dev/null; echo "dag16 rc=$?"
for f in gpurun_out/r2h_bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r = d.get("roofline", {})
    print(sys.argv[1].split("/")[-1], round(d["value"] / 1e6, 3), "M", "frac", r.get("frac"), "frac_step", r.get("frac_step"), "launches", d.get("gpu_launches"), "parity", (d.get("parity") or {}).get("ok"))
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
B="python bench.py --steps 8 --warmup 4 --windows 1 --no-cpu-baseline --no-e2e --no-other-configs --no-parity-gate"
$B $C3 > gpurun_out/plain_r2h_c3.json 2> gpurun_out/plain_r2h_c3.err && \
ncu --set full --clock-control none --import-source on -k regex:k_world_step_pipe -c 2 -f -o gpurun_out/prof_r2h_c3 $B $C3 > gpurun_out/ncu_full_r2h_c3.log 2>&1; echo "c3 full rc=$?"
export FIMO_B200_GENS_HOLD=16
B="python bench.py --steps 32 --warmup 16 --windows 1 --no-cpu-baseline --no-e2e --no-other-configs --no-parity-gate"
$B --dag dag16 --worlds 1024 > gpurun_out/plain_r2h_c5.json 2> gpurun_out/plain_r2h_c5.err && \
ncu --set full --clock-control none --import-source on -k regex:k_world_step_pipe -c 2 -f -o gpurun_out/prof_r2h_c5 $B --dag dag16 --worlds 1024 > gpurun_out/ncu_full_r2h_c5.log 2>&1; echo "c5 full rc=$?"
