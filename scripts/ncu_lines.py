#!/usr/bin/env python
"""Executed warp instructions and stall samples per CUDA source line: joins the SASS table of an ncu report
(source page) with `nvdisasm -g` line info of the cubin extracted from the in-tree .so (same build!).
   python scripts/ncu_lines.py gpurun_out/prof.ncu-rep <kernel-substring> [file-substring] [top_n]"""
import csv
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict

rep, kpat = sys.argv[1], sys.argv[2]
fpat = sys.argv[3] if len(sys.argv) > 3 else "world_step.cu"
top = int(sys.argv[4]) if len(sys.argv) > 4 else 60
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "fimo_b200", "libfimo_worlds_b200.so")

tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=tmp, capture_output=True)
line_of = []  # per instruction (in order) of the first function matching kpat
for cubin in sorted(os.listdir(tmp)):
    if "sm_100a" not in cubin:
        continue
    txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
    cur_fn, cur_line, stack = None, None, []
    for ln in txt.splitlines():
        m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
        if m:
            if line_of:
                break
            cur_fn = m.group(1)
            continue
        if cur_fn is None or kpat not in cur_fn:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', ln)
        if m:
            if "inlined at" in m.group(3):
                m2 = re.search(r'inlined at "([^"]+)", line (\d+)', m.group(3))
                cur_line = (m.group(1), int(m.group(2)), (m2.group(1), int(m2.group(2))) if m2 else None)
            else:
                cur_line = (m.group(1), int(m.group(2)), None)
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            line_of.append((int(m.group(1), 16), cur_line, m.group(2)))
    if line_of:
        break

raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
tables, cur = [], []
for ln in raw.splitlines():
    if ln.startswith('"Address"'):
        if cur:
            tables.append(cur)
        cur = [ln]
    elif cur and not ln.startswith('"Function Name"') and not ln.startswith('"File Path"'):
        cur.append(ln)
if cur:
    tables.append(cur)
rows = list(csv.reader(tables[0]))
hdr = rows[0]
iI, iSamp, iS = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Source")
body = [r for r in rows[1:] if len(r) > max(iI, iSamp)]
if len(body) != len(line_of):
    print(f"warning: {len(body)} SASS rows in the report vs {len(line_of)} in the cubin (different build?)")
agg = defaultdict(lambda: [0, 0])
tot_i = tot_s = 0
for r, (_, li, _) in zip(body, line_of):
    try:
        n, s = int(float(r[iI] or 0)), int(float(r[iSamp] or 0))
    except ValueError:
        continue
    key = "?"
    if li:
        f, l, outer = li
        # attribute to the line in the file of interest (outermost inlining site if the code is from a header)
        if fpat in f or not outer:
            key = f"{os.path.basename(f)}:{l}"
        else:
            key = f"{os.path.basename(outer[0])}:{outer[1]} <- {os.path.basename(f)}:{l}"
    agg[key][0] += n
    agg[key][1] += s
    tot_i += n
    tot_s += s
print(f"total warp instructions {tot_i}, samples {tot_s}")
src_cache = {}
by = 1 if os.environ.get("BY_SAMPLES") else 0
for key, (n, s) in sorted(agg.items(), key=lambda kv: -kv[1][by])[:top]:
    print(f"  {n:10d} {100.0 * n / max(tot_i, 1):5.1f}%  samples {100.0 * s / max(tot_s, 1):5.1f}%  {key}")
