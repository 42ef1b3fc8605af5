#!/usr/bin/env python
"""Sum an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel.  usage: launch_shares.py x.csv"""
import csv
import sys
from collections import defaultdict

rows = list(csv.reader(open(sys.argv[1])))
for i, r in enumerate(rows):
    if "Kernel Name" in r:
        hdr, start = r, i + 1
        break
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
t, c = defaultdict(float), defaultdict(int)
for r in rows[start:]:
    if len(r) <= vi:
        continue
    v = float(r[vi].replace(",", ""))
    v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[ui], 1e-6)
    name = r[ki].split("(")[0][:56]
    t[name] += v
    c[name] += 1
tot = sum(t.values())
for k, v in sorted(t.items(), key=lambda kv: -kv[1])[:14]:
    print(f"{k:58s} {c[k]:4d} launches {v:9.3f} ms {100 * v / tot:5.1f}%")
