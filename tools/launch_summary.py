#!/usr/bin/env python
"""Sums an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel.
usage: python tools/launch_summary.py launches.csv "<header comment>" > profiles/x_launch_summary.txt"""
import csv
import sys
from collections import defaultdict

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = next(r for r in rows if "Kernel Name" in r)
ix = {h: i for i, h in enumerate(hdr)}
tot, cnt = defaultdict(float), defaultdict(int)
for r in rows:
    if r is hdr or r[ix["Metric Name"]] != "gpu__time_duration.sum":
        continue
    v = float(r[ix["Metric Value"]].replace(",", ""))
    u = r[ix["Metric Unit"]]
    ms = v * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0, "second": 1e3}[u]
    tot[r[ix["Kernel Name"]]] += ms
    cnt[r[ix["Kernel Name"]]] += 1
if len(sys.argv) > 2:
    print("# " + sys.argv[2])
total = sum(tot.values())
for k in sorted(tot, key=lambda k: -tot[k]):
    print(f"{tot[k]:10.3f} ms {100 * tot[k] / total:6.2f}% x{cnt[k]:4d}  {k[:130]}")
