#!/usr/bin/env python3
"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list.
usage: python tools/launch_summary.py launches.csv [summary.csv]  (share = share of the time of this library's kernels)"""
import csv
import sys
from collections import OrderedDict

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 14 and r[0].isdigit() and r[12] == "gpu__time_duration.sum"]
agg = OrderedDict()
for r in rows:
    name = r[4]
    ours = "qeb::" in name
    key = ((name if ours else "torch/cublas:" + name)[:70], r[7], r[8])
    unit, val = r[13], float(r[14].replace(",", ""))
    us = val / 1e3 if unit in ("ns", "nsecond") else val * 1e3 if unit in ("ms", "msecond") else val
    a = agg.setdefault(key, [0, 0.0, ours])
    a[0] += 1
    a[1] += us
tot = sum(a[1] for a in agg.values() if a[2])
out = [["kernel", "block", "grid", "launches", "total_us", "avg_us", "share_of_our_kernels"]]
for (name, blk, grid), (n, us, ours) in agg.items():
    out.append([name, blk, grid, n, round(us, 1), round(us / n, 2), round(us / tot, 4) if ours else ""])
w = csv.writer(open(sys.argv[2], "w") if len(sys.argv) > 2 else sys.stdout)
w.writerows(out)
