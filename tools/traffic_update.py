#!/usr/bin/env python3
"""Rewrites profiles/traffic.json from `ncu --set full` summaries (tools/ncu_summary.py output):
   python tools/traffic_update.py --vloc profiles/X_vloc_ncu_full.csv --vloc-boxes 20 --gemm profiles/X_gemm_ncu_full.csv --gemm-bands 800
Per kernel the LAST launch of the capture is used (dram__bytes_read.sum + dram__bytes_write.sum, Mbyte columns)."""
import argparse
import csv
import json
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="ausurf112")
ap.add_argument("--vloc")
ap.add_argument("--vloc-boxes", type=int, default=20)
ap.add_argument("--gemm")
ap.add_argument("--gemm-bands", type=int, default=800)
a = ap.parse_args()
commit = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
path = os.path.join(ROOT, "profiles", "traffic.json")
tj = json.load(open(path))
tj["_doc"] = ("DRAM traffic (dram__bytes_read.sum + dram__bytes_write.sum per launch) of the bench's kernels, read by bench.py for "
              "roofline*.traffic. Written by tools/traffic_update.py from `ncu --set full` summaries under profiles/; each entry names "
              "the capture file and the commit whose kernels were captured.")
w = tj.setdefault(a.workload, {})


def last_per_kernel(fn):
    rows = list(csv.reader(open(fn)))
    h, units = rows[0], rows[1]
    bscale = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}
    tscale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}
    out = {}
    for r in rows[2:]:
        name = r[h.index("Kernel Name")]
        key = "zinv" if "zinv" in name else "zfwd" if "zfwd" in name else "plane" if "plane" in name else "calbec" if "calbec" in name \
            else "addvus" if "addvus" in name else name[:24]
        ir, iw, it = h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum"), h.index("gpu__time_duration.sum")
        mb = float(r[ir]) * bscale[units[ir]] + float(r[iw]) * bscale[units[iw]]  # MB
        out[key] = (mb, float(r[it]) * tscale[units[it]])                             # us
    return out


if a.vloc:
    k = last_per_kernel(a.vloc)
    tot = sum(v[0] for v in k.values()) * 1e6
    w["vloc_bytes_per_box_plane"] = tot / a.vloc_boxes
    w["vloc_source"] = (f"{os.path.relpath(a.vloc, ROOT)} (commit {commit}): " + ", ".join(f"{n} {v[0]:.1f} MB / {v[1]:.0f} us" for n, v in k.items())
                        + f" per batch of {a.vloc_boxes} boxes")
if a.gemm:
    k = last_per_kernel(a.gemm)
    tot = sum(v[0] for v in k.values()) * 1e6
    w["gemm_bytes_per_band"] = tot / a.gemm_bands
    w["gemm_source"] = (f"{os.path.relpath(a.gemm, ROOT)} (commit {commit}): " + ", ".join(f"{n} {v[0] / 1e3:.2f} GB / {v[1] / 1e3:.1f} ms" for n, v in k.items())
                        + f" per launch at m={a.gemm_bands}")
json.dump(tj, open(path, "w"), indent=1)
print(json.dumps(w, indent=1))
