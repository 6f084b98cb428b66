#!/usr/bin/env python
"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list.
usage: launch_summary.py launches.csv > summary.csv"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if r]
start = next(i for i, r in enumerate(rows) if r[0] == "ID")
hdr = rows[start]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot, cnt = collections.Counter(), collections.Counter()
for r in rows[start + 1:]:
    if len(r) <= vi:
        continue
    v = float(r[vi].replace(",", ""))
    scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(r[ui], 1.0)
    tot[r[ki]] += v * scale
    cnt[r[ki]] += 1
allt = sum(tot.values())
print("kernel,launches,total_us,share")
for k, t in tot.most_common():
    print(f"\"{k[:90]}\",{cnt[k]},{t:.1f},{t / allt:.3f}")
