#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum ... --csv` launch list: min..max time per (kernel, grid)."""
import collections, csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]
ki, vi, gi, mi = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size"), hdr.index("Metric Name")
agg = collections.OrderedDict()
for r in rows[1:]:
    if r[mi] != "gpu__time_duration.sum":
        continue
    agg.setdefault(r[ki][:52] + " g=" + r[gi], []).append(float(r[vi].replace(",", "")))
tot = 0
for k, t in agg.items():
    print(f"{min(t)/1e3:8.1f}..{max(t)/1e3:8.1f} us n={len(t):2d} {k}")
