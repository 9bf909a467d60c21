#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel: launches, total and mean time,
share of the listed window.  usage: python tools/launch_shares.py launches.csv > profiles/<name>.md"""
import csv
import sys
from collections import OrderedDict

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10 and r[0] != "ID" and r[0].isdigit()]
agg = OrderedDict()
for r in rows:
    name = r[4].split("(")[0].replace("void ", "").replace("zv::", "")
    ns = float(r[-1])
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += ns
tot = sum(v[1] for v in agg.values())
print(f"{len(rows)} launches, {tot / 1e6:.3f} ms (ncu serialises launches and runs them cold: compare SHARES, not absolute times)\n")
print("| kernel | launches | total ms | mean us | share |")
print("|---|---|---|---|---|")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"| {k} | {v[0]} | {v[1] / 1e6:.3f} | {v[1] / v[0] / 1e3:.1f} | {100 * v[1] / tot:.1f} % |")
