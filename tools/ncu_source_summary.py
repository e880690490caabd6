#!/usr/bin/env python
"""Aggregate an ncu `--page source --print-source cuda,sass --csv` dump by CUDA source line:
instructions executed, stall samples and shared-memory wavefronts per line.
usage: ncu -i X.ncu-rep --page source --print-source cuda,sass --csv | python tools/ncu_source_summary.py [top]"""
import csv
import sys
from collections import defaultdict

top = int(sys.argv[1]) if len(sys.argv) > 1 else 40
rows = list(csv.reader(sys.stdin))
cur_file = None
hdr = None
agg = defaultdict(lambda: defaultdict(float))
src_text = {}
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        continue
    if r[0] == "Function Name" or hdr is None:
        continue
    d = dict(zip(hdr, r))
    try:
        line = int(r[0])
    except ValueError:
        continue
    key = (cur_file, line)
    src_text.setdefault(key, r[1].strip()[:90])
    def f(name):
        try:
            return float(d.get(name, "0") or 0)
        except ValueError:
            return 0.0
    # rows with an Address are SASS rows; the CUDA row aggregates already, so only take SASS rows
    if r[2] == "":
        continue
    a = agg[key]
    a["inst"] += f("Instructions Executed")
    a["samples"] += f("# Samples")
    a["smem_wave"] += f("L1 Wavefronts Shared")
    a["smem_excess"] += f("L1 Wavefronts Shared Excessive")
    for k in ("stall_long_sb", "stall_short_sb", "stall_barrier", "stall_wait", "stall_mio", "stall_math", "stall_not_selected", "stall_branch_resolving", "stall_no_inst", "stall_lg"):
        a[k] += f(k)
tot_inst = sum(a["inst"] for a in agg.values()) or 1
tot_samp = sum(a["samples"] for a in agg.values()) or 1
print(f"total warp-instructions {tot_inst:.3e}, samples {tot_samp:.0f}")
print(f"{'file:line':28s} {'inst%':>6s} {'samp%':>6s} {'long':>5s} {'short':>5s} {'bar':>5s} {'wait':>5s} {'mio':>5s} {'smemW':>9s} {'excess':>8s}  source")
for key, a in sorted(agg.items(), key=lambda kv: -kv[1]["samples"])[:top]:
    s = a["samples"] or 1
    print(f"{key[0][:20]}:{key[1]:<6d} {100*a['inst']/tot_inst:6.2f} {100*a['samples']/tot_samp:6.2f} "
          f"{100*a['stall_long_sb']/s:5.0f} {100*a['stall_short_sb']/s:5.0f} {100*a['stall_barrier']/s:5.0f} {100*a['stall_wait']/s:5.0f} {100*a['stall_mio']/s:5.0f} "
          f"{a['smem_wave']:9.3g} {a['smem_excess']:8.3g}  {src_text[key]}")
