#!/usr/bin/env python3
"""Parameter sweep of the V_loc section on one GPU (batch size, warps per CTA). Prints us/band."""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import qe_b200  # noqa: E402
from bench import workload  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="ausurf112")
ap.add_argument("--bands", type=int, default=128)
ap.add_argument("--batches", default="4")
ap.add_argument("--env", default="", help="semicolon separated settings, e.g. 'QEB_NWX=5;QEB_NWX=8,QEB_NWY=4'")
a = ap.parse_args()
d, c, nh = workload(a.workload)
n = lda = d.npw
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
m = a.bands
psi = torch.randn((m, lda, 2), generator=g, device=dev, dtype=torch.float64) / np.sqrt(2 * n)
hpsi = torch.zeros_like(psi)
rng = np.random.default_rng(0)
vrs = rng.uniform(-1, 0, d.nnr)
s = torch.cuda.Stream()
torch.cuda.set_stream(s)
for batch in [int(x) for x in a.batches.split(",")]:
    for setting in (a.env.split(";") if a.env else [""]):
        keys = []
        for kv in setting.split(","):
            if "=" in kv:
                k, v = kv.split("=", 1)
                os.environ[k] = v.replace("+", ",")  # radix lists are written 5+5+5 inside a setting
                keys.append(k)
        # the context is created per setting: the plan overrides (QEB_[XYZ]PLAN) are read by qeb_set_grid
        h = qe_b200.HPsi(stream=s.cuda_stream, batch=batch)
        h.set_grid(d.nr1, d.nr2, d.nr3, d.nl, d.nlm if d.gamma_only else None, d.gamma_only)
        h.set_kpoint(n, lda, None if d.gamma_only else d.igk, d.g2kin, None, 0, d.gstart)
        h.set_vloc(vrs)
        h.set_async(True)
        f = h.vloc_psi_gamma if d.gamma_only else h.vloc_psi_k
        for _ in range(2):
            f(lda, n, m, psi, None, hpsi)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        reps = 3
        for _ in range(reps):
            f(lda, n, m, psi, None, hpsi)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / reps / m
        st = h.stats()
        print(f"batch={batch:3d} {setting:40s} {us:8.2f} us/band  {st['vloc_bytes_per_band'] / us / 1e3:8.1f} GB/s", flush=True)
        h.close()
        for k in keys:
            del os.environ[k]
