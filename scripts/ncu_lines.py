#!/usr/bin/env python3
"""Aggregate an `ncu --page source --print-source cuda,sass --csv` export per CUDA source line:
share of executed warp instructions, share of stall samples and the dominant stall reasons."""
import csv
import sys
from collections import defaultdict

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(open(path)))
agg = defaultdict(lambda: {"inst": 0, "samp": 0, "stalls": defaultdict(int), "src": ""})
cur_file, hdr = None, None
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if r[0] == "Function Name":
        continue
    if r[0] == "Line No":
        hdr = r
        ii, si = hdr.index("Instructions Executed"), hdr.index("# Samples")
        stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
        continue
    if hdr is None or len(r) < len(hdr):
        continue
    try:
        inst, samp = int(r[ii]), int(r[si])
    except ValueError:
        continue
    if r[0] != "":
        last_line = int(r[0])
        last_src = r[1].strip()[:100]
        continue  # the CUDA line row repeats the sum of its SASS rows
    key = (cur_file, last_line)
    a = agg[key]
    a["inst"] += inst
    a["samp"] += samp
    a["src"] = last_src
    for i, h in stall_cols:
        try:
            a["stalls"][h] += int(r[i])
        except ValueError:
            pass
ti = sum(a["inst"] for a in agg.values())
ts = sum(a["samp"] for a in agg.values())
print(f"total warp instructions {ti}, samples {ts}")
glob = defaultdict(int)
for a in agg.values():
    for h, v in a["stalls"].items():
        glob[h] += v
print("stall mix:", ", ".join(f"{h[6:]} {100 * v / max(ts, 1):.1f}%" for h, v in sorted(glob.items(), key=lambda kv: -kv[1])[:8]))
for title, keyf in (("by instructions", lambda kv: -kv[1]["inst"]), ("by stall samples", lambda kv: -kv[1]["samp"])):
    print(f"--- {title}")
    for (f, line), a in sorted(agg.items(), key=keyf)[:top]:
        st = ", ".join(f"{h[6:]} {v}" for h, v in sorted(a["stalls"].items(), key=lambda kv: -kv[1])[:3] if v)
        print(f"{100 * a['inst'] / ti:5.1f}% inst {100 * a['samp'] / max(ts, 1):5.1f}% samp  {f}:{line}  {a['src']}   [{st}]")
