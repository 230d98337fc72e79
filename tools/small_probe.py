#!/usr/bin/env python3
"""Launch-bound workloads: us per h_psi call of one Si-2 k-point, eager vs graph replay, and of 16 k-points per call."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import qe_b200  # noqa: E402
from bench import workload  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "si2"
nk = int(sys.argv[2]) if len(sys.argv) > 2 else 16
dev = torch.device("cuda", 0)
s = torch.cuda.Stream()
torch.cuda.set_stream(s)
ctxs, psis, outs, ns = [], [], [], []
m = None
for k in range(nk):
    d, c, nh = workload(name, seed=k)
    m = c["nbnd"]
    n = d.npw
    nkb = int(nh.sum())
    g = torch.Generator(device=dev).manual_seed(k)
    psi = torch.randn((m, n, 2), generator=g, device=dev, dtype=torch.float64)
    vkb = torch.randn((nkb, n, 2), generator=g, device=dev, dtype=torch.float64) / np.sqrt(n)
    h = qe_b200.HPsi(stream=s.cuda_stream) if k == 0 else qe_b200.HPsi()
    h.set_grid(d.nr1, d.nr2, d.nr3, d.nl, None, False)
    h.set_kpoint(n, n, d.igk, d.g2kin, vkb, nkb, d.gstart)
    h.set_vloc(np.random.default_rng(0).uniform(-1, 0, d.nnr))
    nhm = int(nh.max())
    deeq = np.random.default_rng(1).uniform(-1, 1, (len(nh), nhm, nhm))
    h.set_nonlocal(nh, np.concatenate([[0], np.cumsum(nh)[:-1]]).astype(np.int32), 0.5 * (deeq + deeq.transpose(0, 2, 1)))
    h.set_async(True)
    ctxs.append(h)
    psis.append(psi)
    outs.append(torch.zeros_like(psi))
    ns.append(n)


def timeit(fn, reps=200):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e6


for mode in ("0", "1"):
    os.environ["QEB_GRAPH"] = mode
    us1 = timeit(lambda: ctxs[0].h_psi(ns[0], ns[0], m, psis[0], outs[0]))
    usk = timeit(lambda: qe_b200.h_psi_multi(ctxs, ns, ns, [m] * nk, psis, outs), reps=50)
    print(f'{{"probe":"small","workload":"{name}","graph":{mode},"us_per_call_1k":{us1:.1f},"band_apps_per_s_1k":{m / us1 * 1e6:.0f},'
          f'"us_per_call_{nk}k":{usk:.1f},"band_apps_per_s_{nk}k":{nk * m / usk * 1e6:.0f}}}', flush=True)
