#!/usr/bin/env python3
"""Executed warp instructions and stall samples per source FUNCTION of one kernel, from an
`ncu --page source --print-source cuda,sass --csv` export.  SASS of inlined helpers (shuffles, __ldg, popc helpers in
other headers) is charged to the enclosing function of the main file by address order: every SASS row belongs to the
last preceding row (by address) that maps to a line of MAIN_FILE.

usage: ncu_regions.py SRC.csv MAIN_FILE.cuh name:first_line[,name:first_line...]"""
import csv
import sys
from collections import defaultdict

path, main = sys.argv[1], sys.argv[2]
regions = sorted(((int(x.split(":")[1]), x.split(":")[0]) for x in sys.argv[3].split(",")))
rows = list(csv.reader(open(path)))
sass = []
cur_file = hdr = None
last_line = 0
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        ii, si = hdr.index("Instructions Executed"), hdr.index("# Samples")
        continue
    if r[0] == "Function Name" or hdr is None or len(r) < len(hdr):
        continue
    if r[0] != "":
        last_line = int(r[0])
        continue
    try:
        sass.append((int(r[2], 16), int(r[ii]), int(r[si]), cur_file, last_line))
    except ValueError:
        pass
sass.sort()
agg = defaultdict(lambda: [0, 0])
ctx = 0
for addr, inst, samp, f, line in sass:
    if f == main:
        ctx = line
    name = "?"
    for first, nm in regions:
        if ctx >= first:
            name = nm
    agg[name][0] += inst
    agg[name][1] += samp
ti = sum(v[0] for v in agg.values())
ts = sum(v[1] for v in agg.values())
for nm, (i, s) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    print(f"{100 * i / ti:6.2f}% inst {100 * s / max(ts, 1):6.2f}% samples  {nm}")
