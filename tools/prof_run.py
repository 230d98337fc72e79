#!/usr/bin/env python3
"""Small fixed workload for ncu: AUSURF112 tables, m bands, `calls` h_psi calls on device buffers."""
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
ap.add_argument("--bands", type=int, default=64)
ap.add_argument("--calls", type=int, default=2)
ap.add_argument("--batch", type=int, default=0)
ap.add_argument("--mode", default="serial", choices=["serial", "rg", "nc"])
a = ap.parse_args()
d, c, nh = workload(a.workload)
n = lda = d.npw
nkb = int(nh.sum())
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
psi = torch.randn((a.bands, lda, 2), generator=g, device=dev, dtype=torch.float64) / np.sqrt(2 * n)
hpsi = torch.zeros_like(psi)
vkb = torch.randn((nkb, lda, 2), generator=g, device=dev, dtype=torch.float64) / np.sqrt(n)
rng = np.random.default_rng(0)
nhm = int(nh.max())
deeq = rng.uniform(-1, 1, (len(nh), nhm, nhm))
ofs = np.concatenate([[0], np.cumsum(nh)[:-1]]).astype(np.int32)
h = qe_b200.HPsi(batch=a.batch or None)
if a.mode == "rg":  # the slab-distributed code path on one rank (exchange = device copy)
    d.distribute(1)
    rk = d.ranks[0]
    h.set_grid_dist(d.nr1, d.nr2, d.nr3, 1, 0, d.nr3p, d.nsw, d.iss, d.ismap, rk.nl, rk.nlm if d.gamma_only else None,
                    d.gamma_only)
    h.set_kpoint(n, lda, None if d.gamma_only else rk.igk, rk.g2kin, vkb, nkb, rk.gstart)
else:
    h.set_grid(d.nr1, d.nr2, d.nr3, d.nl, d.nlm if d.gamma_only else None, d.gamma_only)
    h.set_kpoint(n, lda, None if d.gamma_only else d.igk, d.g2kin, vkb, nkb, d.gstart)
h.set_vloc(rng.uniform(-1, 0, d.nnr))
h.set_nonlocal(nh, ofs, deeq)
if a.mode == "nc":  # noncollinear V_loc: the bands as bands/2 spinors through the 2x2 potential
    h.set_vloc_nc(rng.uniform(-1, 1, (4, d.nnr)))
for _ in range(a.calls):
    if a.mode == "nc":
        h.vloc_psi_nc(lda, n, a.bands // 2, psi, None, hpsi)
    else:
        h.h_psi(lda, n, a.bands, psi, hpsi)
torch.cuda.synchronize()
print("ok", h.stats())
