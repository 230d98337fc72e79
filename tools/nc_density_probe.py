#!/usr/bin/env python3
"""Timing of the density loops at AUSURF112 size on one GPU (device-resident inputs, CUDA events on the library's stream):
qeb_sum_band_rho on 800 columns, qeb_sum_band_rho_nc on the same columns read as 400 spinors with and without
magnetisation. One JSON line per item."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import qe_b200  # noqa: E402
from oracle import descriptor as OD  # noqa: E402  (input generation only: grid maps of the workload)

d = OD.config_descriptor("ausurf112")
n = lda = d.npw
m = 800
dev = torch.device("cuda", 0)
s = torch.cuda.Stream()
torch.cuda.set_stream(s)
h = qe_b200.HPsi(stream=s.cuda_stream)
h.set_async(True)
h.set_grid(d.nr1, d.nr2, d.nr3, d.nl)
h.set_kpoint(n, lda, d.igk, d.g2kin, None, 0, d.gstart)


def timed(fn, reps=3, warm=1):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


g = torch.Generator(device=dev).manual_seed(1)
psi = torch.randn((m, lda, 2), generator=g, device=dev, dtype=torch.float64) / np.sqrt(2 * n)
w = torch.full((m,), 2.0e-3, device=dev, dtype=torch.float64)
rho = torch.zeros((4, d.nnr), device=dev, dtype=torch.float64)
ms = timed(lambda: h.sum_band_rho(lda, n, m, psi, w, rho[0]))
print(json.dumps({"item": "sum_band_rho, 800 columns", "ms": ms, "columns_per_s": m / (ms * 1e-3)}), flush=True)
ms = timed(lambda: h.sum_band_rho_nc(lda, n, m // 2, psi, w, rho[0], domag=False))
print(json.dumps({"item": "sum_band_rho_nc, 400 spinors, domag=0", "ms": ms, "spinors_per_s": (m // 2) / (ms * 1e-3)}), flush=True)
ms = timed(lambda: h.sum_band_rho_nc(lda, n, m // 2, psi, w, rho, domag=True))
print(json.dumps({"item": "sum_band_rho_nc, 400 spinors, domag=1 (four scalar densities + fold)", "ms": ms,
                  "spinors_per_s": (m // 2) / (ms * 1e-3)}), flush=True)
h.close()
