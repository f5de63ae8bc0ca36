#!/usr/bin/env python
"""Per-source-line instruction and stall totals from an ncu report's source page (needs -lineinfo):
   python scripts/ncu_source.py gpurun_out/prof.ncu-rep [top_n]"""
import csv
import subprocess
import sys
from collections import defaultdict

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
lines = raw.splitlines()
# the page holds one table per kernel launch; take the first
tables, cur = [], []
for ln in lines:
    if ln.startswith('"Function Name"') or ln.startswith('"File Path"'):
        continue
    if ln.startswith('"Address"') or ln.startswith('"Line No"'):
        if cur:
            tables.append(cur)
        cur = [ln]
    elif cur:
        cur.append(ln)
if cur:
    tables.append(cur)
rows = list(csv.reader(tables[0]))
hdr = rows[0]
iS, iI, iSamp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
tot_i = tot_s = 0
ops = defaultdict(lambda: [0, 0])
out = []
for r in rows[1:]:
    if len(r) <= max(iS, iI, iSamp):
        continue
    try:
        n, s = int(float(r[iI] or 0)), int(float(r[iSamp] or 0))
    except ValueError:
        continue
    tot_i += n
    tot_s += s
    op = r[iS].split()[0] if r[iS].split() else "?"
    if op.startswith("@"):
        op = r[iS].split()[1]
    ops[op.split(".")[0]][0] += n
    ops[op.split(".")[0]][1] += s
    out.append((n, s, r[iS][:90]))
print(f"total warp instructions {tot_i}, samples {tot_s}")
print("by opcode (instr, samples):")
for op, (n, s) in sorted(ops.items(), key=lambda kv: -kv[1][0])[:25]:
    print(f"  {op:12s} {n:12d} {100.0 * n / max(tot_i, 1):5.1f}%   samples {100.0 * s / max(tot_s, 1):5.1f}%")
print("hottest SASS by samples:")
for n, s, src in sorted(out, key=lambda t: -t[1])[:top]:
    print(f"  {s:7d} {n:10d}  {src}")
