#!/usr/bin/env python
"""Median gpu__time_duration per kernel from an `ncu --csv --metrics gpu__time_duration.sum` log (serialised, cold-cache launches)."""
import csv, sys, statistics, re
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10 and r[0].isdigit()]
d = {}
for r in rows:
    name = re.sub(r"\(.*", "", r[4]).replace("void ", "").replace("sdrx::", "")
    d.setdefault(name, []).append(float(r[-1]) / 1e3)
for k, v in sorted(d.items(), key=lambda kv: -statistics.median(kv[1])):
    print(f"{k:40s} n {len(v):3d}  median {statistics.median(v):7.1f} us  min {min(v):7.1f}  max {max(v):7.1f}")
