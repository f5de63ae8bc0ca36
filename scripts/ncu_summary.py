#!/usr/bin/env python
"""Condenses an ncu report (read here, no GPU) into the few numbers the roofline needs:
   python scripts/ncu_summary.py gpurun_out/prof.ncu-rep [kernel-substring] > profiles/<name>.txt"""
import csv
import subprocess
import sys

rep = sys.argv[1]
pat = sys.argv[2] if len(sys.argv) > 2 else ""
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second",
        "dram__bytes_write.sum.per_second", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "sm__cycles_elapsed.max",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_sectors_srcunit_tex_op_write.sum"]
for r in rows[2:]:
    d = dict(zip(hdr, r))
    if pat and pat not in d.get("Kernel Name", ""):
        continue
    print("kernel:", d.get("Kernel Name"), "id", d.get("ID"))
    for k in KEYS:
        if k in d:
            print(f"  {k:70s} {d[k]:>16s} {units[hdr.index(k)]}")
    try:
        print(f"  traffic (dram read+write)                                             "
              f"{float(d['dram__bytes_read.sum']) + float(d['dram__bytes_write.sum']):16.3f} {units[hdr.index('dram__bytes_read.sum')]}")
    except Exception:
        pass
    ps = []
    for k, v in d.items():
        if k.startswith("smsp__pcsamp_warps_issue_stalled") and not k.endswith("not_issued"):
            try:
                ps.append((float(v.replace(",", "")), k))
            except ValueError:
                pass
    tot = sum(p for p, _ in ps) or 1.0
    print("  warp stall samples:")
    for p, k in sorted(ps, reverse=True)[:8]:
        print(f"    {100 * p / tot:5.1f}%  {k.replace('smsp__pcsamp_warps_issue_stalled_', '')}")
