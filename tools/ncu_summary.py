#!/usr/bin/env python
"""Condense `ncu -i X.ncu-rep --page raw --csv` into a small per-kernel table (markdown) for profiles/.

usage: ncu -i rep.ncu-rep --page raw --csv > raw.csv; python tools/ncu_summary.py raw.csv > profiles/<name>.md
"""
import csv
import sys
from collections import OrderedDict

COLS = OrderedDict([
    ("gpu__time_duration.sum", "us"),
    ("dram__bytes_read.sum", "dram_rd"),
    ("dram__bytes_write.sum", "dram_wr"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"),
    ("sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tmem/tensor%"),
    ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed", "xu%"),
    ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_elapsed", "fma%"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_elapsed", "alu%"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
    ("launch__registers_per_thread", "regs"),
    ("launch__shared_mem_per_block_dynamic", "dsmem"),
])


def main(path):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    cols = [(k, v) for k, v in COLS.items() if k in ix]
    print("| kernel | grid | block | " + " | ".join(v for _, v in cols) + " |")
    print("|---|---|---|" + "---|" * len(cols))
    for r in data:
        name = r[ix["Kernel Name"]].split("(")[0].replace("void ", "")
        vals = []
        for k, _ in cols:
            v, u = r[ix[k]], units[ix[k]]
            try:
                f = float(v)
                v = f"{f:.1f}" if abs(f) < 1e4 else f"{f:.3g}"
            except ValueError:
                pass
            vals.append(f"{v} {u}".strip() if u not in ("%", "") and "byte" in u else v)
        print(f"| {name} | {r[ix['Grid Size']]} | {r[ix['Block Size']]} | " + " | ".join(vals) + " |")


if __name__ == "__main__":
    main(sys.argv[1])
