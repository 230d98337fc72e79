#!/usr/bin/env python3
"""Per-kernel SASS evidence of libqeb200.so: counts of the instructions that identify the hardware paths used
(DMMA = FP64 tensor cores, UBLKCP = bulk async copy (TMA engine), SYNCS = mbarrier, LDGSTS = cp.async, DFMA/DADD/DMUL =
FP64 vector pipe). usage: python tools/sass_summary.py [lib.so] [out.csv]"""
import collections
import csv
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "q-e_b200/libqeb200.so"
out = sys.argv[2] if len(sys.argv) > 2 else None
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
want = ["DMMA", "UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "LDGSTS", "DFMA", "DADD", "DMUL", "LDS", "STS", "LDG", "STG", "BAR",
        "LDL", "STL", "RED", "ATOM"]
rows = []
name, cnt = None, None
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        if name:
            rows.append((name, cnt))
        name, cnt = m.group(1), collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and name:
        op = m.group(1)
        cnt["total"] += 1
        for w in want:
            if op == w or op.startswith(w + "."):
                cnt[w] += 1
if name:
    rows.append((name, cnt))
dem = subprocess.run(["cu++filt"] + [r[0] for r in rows], capture_output=True, text=True).stdout.splitlines()
table = [["kernel", "total"] + want]
for (n, c), d in zip(rows, dem if len(dem) == len(rows) else [r[0] for r in rows]):
    short = re.sub(r"\((int|bool)\)", "", d).replace("qeb::", "").replace("void ", "")
    short = re.sub(r"\(.*", "", short)
    table.append([short, c["total"]] + [c[w] for w in want])
table[1:] = sorted(table[1:], key=lambda r: r[0])
if out:
    csv.writer(open(out, "w")).writerows(table)
for r in table:
    if r[0] == "kernel" or any(k in r[0] for k in ("plane2", "calbec3m", "addvus3m", "zinv_f", "zfwd_f")):
        print(",".join(str(x) for x in r[:12]))
