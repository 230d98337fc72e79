#!/usr/bin/env python3
"""Summarise an `ncu --page source --csv` export: total stall samples by reason and the hottest SASS lines.
usage: ncu -i rep --page source --csv -k regex:NAME > f.csv; python tools/ncu_stalls.py f.csv [top]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hi = next(i for i, r in enumerate(rows[:40]) if "Source" in r and "# Samples" in r)
hdr = rows[hi]
col = {n: i for i, n in enumerate(hdr)}
stall_cols = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
body = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
tot = {n: 0 for n in stall_cols}
for r in body:
    for n in stall_cols:
        try:
            tot[n] += int(r[col[n]])
        except ValueError:
            pass
alls = sum(tot.values())
print("kernel:", rows[0][1][:100])
print("stall samples:", alls)
for n, v in sorted(tot.items(), key=lambda kv: -kv[1]):
    if v:
        print(f"  {n:28s} {v:8d} {100.0 * v / max(alls, 1):5.1f}%")


def samples(r):
    try:
        return int(r[col["# Samples"]])
    except ValueError:
        return 0


print("hottest lines:")
for r in sorted(body, key=samples, reverse=True)[:top]:
    reasons = sorted(((int(r[col[n]] or 0), n) for n in stall_cols), reverse=True)[:2]
    print(f"  {samples(r):7d}  {r[col['Source']][:90]:90s} {reasons}")
