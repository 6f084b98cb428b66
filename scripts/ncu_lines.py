#!/usr/bin/env python
"""Per-source-line share of executed warp instructions for the first kernel of an .ncu-rep
(needs -lineinfo and --import-source on).  usage: ncu_lines.py rep [kernel-index] [top]"""
import collections
import csv
import subprocess
import sys

rep = sys.argv[1]
which = int(sys.argv[2]) if len(sys.argv) > 2 else 1
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
res, smp = collections.Counter(), collections.Counter()
cur, fn, first_file = None, 0, None
for r in csv.reader(raw.splitlines()):
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
        if first_file is None:
            first_file = cur
        if cur == first_file:
            fn += 1
        continue
    if fn != which or not r[0].isdigit() or len(r) < 9 or not r[7].isdigit():
        continue
    key = (cur, int(r[0]), r[1].strip()[:72])
    res[key] += int(r[7])
    smp[key] += int(r[6]) if r[6].isdigit() else 0
tot, ts = sum(res.values()), max(sum(smp.values()), 1)
print(f"# {rep}: kernel #{which}, executed warp instructions {tot}")
byfile = collections.Counter()
for (f, _, _), n in res.items():
    byfile[f] += n
print("# by file:", {k: f"{v / tot * 100:.1f}%" for k, v in byfile.items() if v})
for key, n in res.most_common(top):
    print(f"{n / tot * 100:5.1f}% inst  {smp[key] / ts * 100:5.1f}% samples  {key[0]}:{key[1]}  {key[2]}")
