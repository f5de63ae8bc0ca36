#!/bin/bash
# 4 GPUs: does binding each rank to its GPU's NUMA node lift the D2H ceiling of the read-back e2e leg (79 M at N=4 unbound)?
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2m_topo.txt 2>&1
for d in /sys/bus/pci/devices/*; do if [ -f $d/numa_node ] && grep -qi nvidia $d/uevent 2>/dev/null; then echo "$d $(cat $d/numa_node)"; fi; done >> gpurun_out/r2m_topo.txt
lscpu | grep -i "numa\|socket\|^CPU(s)" >> gpurun_out/r2m_topo.txt; nproc >> gpurun_out/r2m_topo.txt
T="timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus 4 --steps 20 --warmup 5"
$T > gpurun_out/r2m_bench_c2_4gpu_numa.json 2> gpurun_out/r2m_bench_c2_4gpu_numa.err; echo "numa rc=$?"
FIMO_B200_NO_NUMA_BIND=1 $T > gpurun_out/r2m_bench_c2_4gpu_nobind.json 2> gpurun_out/r2m_bench_c2_4gpu_nobind.err; echo "nobind rc=$?"
for f in gpurun_out/r2m_bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1].split("/")[-1], "N", d["n_gpus"], round(d["value"] / 1e6, 3), "M", "e2e", round(d["e2e"]["value"] / 1e6, 2), "const", round(d["e2e_constants_only"]["value"] / 1e6, 2), d["config"].get("numa"), d["clocks"])
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
cat gpurun_out/r2m_topo.txt | tail -30
