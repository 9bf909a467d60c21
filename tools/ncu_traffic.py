#!/usr/bin/env python
"""Per-kernel DRAM traffic from `ncu -i X.ncu-rep --page raw --csv`: mean dram__bytes_read/write.sum and duration
per launch, keyed by kernel name.  usage: python tools/ncu_traffic.py raw.csv > profiles/<tag>_traffic.json"""
import csv
import json
import sys
from collections import OrderedDict

rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3}


def val(r, key):
    return float(r[ix[key]].replace(",", "")) * SCALE[units[ix[key]]]


agg = OrderedDict()
for r in data:
    name = r[ix["Kernel Name"]].split("(")[0].replace("void ", "").replace("zv::", "")
    a = agg.setdefault(name, [0, 0.0, 0.0, 0.0])
    a[0] += 1
    a[1] += val(r, "dram__bytes_read.sum")
    a[2] += val(r, "dram__bytes_write.sum")
    a[3] += val(r, "gpu__time_duration.sum")
out = OrderedDict((k, {"captured_launches": v[0], "dram_read_bytes_per_launch": v[1] / v[0],
                       "dram_write_bytes_per_launch": v[2] / v[0], "mean_us": v[3] / v[0]}) for k, v in agg.items())
json.dump(out, sys.stdout, indent=1)
