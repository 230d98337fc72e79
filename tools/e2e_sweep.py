#!/usr/bin/env python3
"""e2e (pinned host psi/hpsi) time of qeb_h_psi at AUSURF112 for several pipeline chunk sizes."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import qe_b200
from bench import workload
d, c, nh = workload("ausurf112")
n = lda = d.npw; m = 800; nkb = int(nh.sum())
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
vkb = torch.randn((nkb, lda, 2), generator=g, device=dev, dtype=torch.float64) / np.sqrt(n)
rng = np.random.default_rng(0)
nhm = int(nh.max()); deeq = rng.uniform(-1, 1, (len(nh), nhm, nhm))
ofs = np.concatenate([[0], np.cumsum(nh)[:-1]]).astype(np.int32)
h = qe_b200.HPsi()
h.set_grid(d.nr1, d.nr2, d.nr3, d.nl); h.set_kpoint(n, lda, d.igk, d.g2kin, vkb, nkb, d.gstart)
h.set_vloc(rng.uniform(-1, 0, d.nnr)); h.set_nonlocal(nh, ofs, deeq)
psi_h = torch.randn((m, lda, 2), dtype=torch.float64).pin_memory(); hpsi_h = torch.empty_like(psi_h).pin_memory()
pn, hn = psi_h.numpy(), hpsi_h.numpy()
for chunk in [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "0,96,128,192,256,400").split(",")]:
    h.set_pipeline(chunk, 128)
    h.h_psi(lda, n, m, pn, hn); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3): h.h_psi(lda, n, m, pn, hn)
    torch.cuda.synchronize()
    print(f"chunk={chunk:4d}  {(time.perf_counter() - t0) / 3 * 1e3:8.2f} ms", flush=True)
h.close()
