#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: total time, share and launch count per kernel.

usage: python scripts/ncu_launch_summary.py LAUNCHES.csv ["command line that produced it"]
"""
import csv
import sys
from collections import defaultdict


def main() -> None:
    path = sys.argv[1]
    rows = []
    with open(path, newline="") as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(unit, 1e-6)
        rows.append((r["Kernel Name"], v))
    tot = sum(v for _, v in rows)
    agg = defaultdict(lambda: [0.0, 0])
    for k, v in rows:
        agg[k][0] += v
        agg[k][1] += 1
    if len(sys.argv) > 2:
        print(sys.argv[2])
    for k, (v, n) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        print(f"{v:12.3f} ms {100 * v / tot:7.3f}%  launches={n:5d}  {k[:110]}")
    print(f"\n{tot:12.3f} ms total over {len(rows)} launches")


if __name__ == "__main__":
    main()
