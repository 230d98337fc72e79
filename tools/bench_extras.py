#!/usr/bin/env python3
"""Timing of the SURVEY 8 "next" rows at AUSURF112 size on one GPU (device-resident inputs, CUDA events on the library's
stream): init_us_2 / g2_kin on the device vs the vkb upload they replace, s_psi, the sum_band density loop and the
Davidson reduced-matrix / basis-update GEMMs. Prints one JSON line per item."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import qe_b200  # noqa: E402
from oracle import descriptor as OD  # noqa: E402  (input generation only: G-vectors, Miller indices, bg)
from oracle import uspp_ref as U  # noqa: E402  (synthetic PP tables only: input generation, not the measured path)

d = OD.config_descriptor("ausurf112")
n = lda = d.npw
nat, m = 112, 800
dev = torch.device("cuda", 0)
s = torch.cuda.Stream()
torch.cuda.set_stream(s)
h = qe_b200.HPsi(stream=s.cuda_stream)
h.set_async(True)


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


def out(**kw):
    print(json.dumps(kw), flush=True)


us = U.synthetic_uspp(1, 2, nbeta_per_l=2)  # nh = 18 per atom -> nkb = 2016
nkb = int(us["nh"][0]) * nat
rng = np.random.default_rng(0)
tau = rng.uniform(-0.5, 0.5, (nat, 3))
ityp = np.ones(nat, dtype=np.int32)
nr_dense = (180, 90, 288)
eig = U.eigts(d.bg, tau, nr_dense)
tpiba = 2.0 * np.pi / d.alat
h.set_grid(d.nr1, d.nr2, d.nr3, d.nl)
h.set_kpoint(n, lda, d.igk, None, None, nkb, d.gstart)
h.set_gvectors(d.g, d.mill)
h.set_beta_tables(us)
h.set_structure(ityp, tau, nr_dense, eig)
ms = timed(lambda: h.g2_kin(d.xk, d.tpiba2))
out(item="g2_kin on device", ms=ms, npw=n)
ms = timed(lambda: h.init_us_2(d.xk, tpiba))
out(item="init_us_2 on device (vkb 16*npwx*nkb bytes written)", ms=ms, nkb=nkb, gb_written=16e-9 * lda * nkb,
    gbs=16e-9 * lda * nkb / (ms * 1e-3))
# what it replaces: uploading a host vkb at every k-point change
vkb_h = torch.empty((nkb, lda, 2), dtype=torch.float64).pin_memory()
vkb_d = torch.empty((nkb, lda, 2), dtype=torch.float64, device=dev)
ms_up = timed(lambda: vkb_d.copy_(vkb_h, non_blocking=True))
out(item="H2D upload of vkb from pinned host memory (the path init_us_2 on the device removes)", ms=ms_up,
    gbs=16e-9 * lda * nkb / (ms_up * 1e-3))
del vkb_h, vkb_d
# h_psi + s_psi on the device-built state
g = torch.Generator(device=dev).manual_seed(1)
psi = torch.randn((m, lda, 2), generator=g, device=dev, dtype=torch.float64) / np.sqrt(2 * n)
hpsi, spsi = torch.empty_like(psi), torch.empty_like(psi)
nh_at = np.full(nat, 18, dtype=np.int32)
ofs = (np.arange(nat) * 18).astype(np.int32)
deeq = rng.uniform(-1, 1, (nat, 18, 18))
qq = rng.uniform(-0.5, 0.5, (nat, 18, 18))
h.set_vloc(rng.uniform(-1, 0, d.nnr))
h.set_nonlocal(nh_at, ofs, deeq)
h.set_qq(qq)
ms_h = timed(lambda: h.h_psi(lda, n, m, psi, hpsi))
out(item="h_psi with device-built vkb/g2kin", ms=ms_h, band_apps_per_s=m / (ms_h * 1e-3))
ms = timed(lambda: h.s_psi(lda, n, m, psi, spsi))
out(item="s_psi (re-uses becp of h_psi)", ms=ms, band_apps_per_s=m / (ms * 1e-3),
    tflops_algorithmic=8.0 * nkb * n * m / (ms * 1e-3) / 1e12)
# sum_band density loop
w = torch.full((m,), 2.0 / 1000.0, device=dev, dtype=torch.float64)
rho = torch.zeros(d.nnr, device=dev, dtype=torch.float64)
ms = timed(lambda: h.sum_band_rho(lda, n, m, psi, w, rho))
out(item="sum_band density loop (scatter + 3-D iFFT + rho += w|psi|^2)", ms=ms, bands_per_s=m / (ms * 1e-3))
# Davidson GEMMs: hc = psi^H hpsi (800 x 800 x npw), psi_new = hpsi vc (npw x 800 x 800)
hc = torch.empty((m, m, 2), device=dev, dtype=torch.float64)
ms = timed(lambda: h.dav_gemm_cn(n, psi, lda, m, hpsi, lda, m, hc, m))
out(item="Davidson reduced matrix psi^H hpsi (ZGEMM C,N 800x800xnpw)", ms=ms, tflops_algorithmic=8.0 * m * m * n / (ms * 1e-3) / 1e12)
vc = torch.randn((m, m, 2), generator=g, device=dev, dtype=torch.float64)
ms = timed(lambda: h.dav_gemm_nn(n, hpsi, lda, m, vc, m, m, spsi, lda))
out(item="Davidson basis update hpsi vc (ZGEMM N,N npwx800x800)", ms=ms, tflops_algorithmic=8.0 * m * m * n / (ms * 1e-3) / 1e12)
e = torch.linspace(-1, 1, m, device=dev, dtype=torch.float64)
h.set_g_psi_diag(np.ascontiguousarray(d.g2kin + 1.0), np.ones(n))
ms = timed(lambda: h.g_psi(lda, n, m, 1, spsi, e))
out(item="g_psi preconditioner", ms=ms, gbs=32e-9 * n * m / (ms * 1e-3))
# noncollinear V_loc with the 2x2 potential: the 800 columns as 400 spinors; the scalar path on the same columns beside it
v4 = torch.randn((4, d.nnr), generator=g, device=dev, dtype=torch.float64)
h.set_vloc_nc(v4)
ms = timed(lambda: h.vloc_psi_nc(lda, n, m // 2, psi, None, hpsi))
ms_k = timed(lambda: h.vloc_psi_k(lda, n, m, psi, None, hpsi))
out(item="vloc_psi_nc (400 spinors = 800 component FFT pairs, spinor x pass)", ms=ms, us_per_component=1e3 * ms / m,
    scalar_vloc_psi_k_same_columns_ms=ms_k)
# whole noncollinear h_psi (kinetic + vloc_psi_nc + calbec_nc + add_vuspsi_nc) and s_psi_nc on the same 400 spinors
dnc = torch.randn((4, nat, 18, 18, 2), generator=g, device=dev, dtype=torch.float64)
h.set_nonlocal_nc(dnc)
ms = timed(lambda: h.h_psi_nc(lda, n, m // 2, psi, hpsi))
out(item="h_psi_nc (400 spinors: 2x2 potential, deeq_nc)", ms=ms, spinor_apps_per_s=(m // 2) / (ms * 1e-3),
    scalar_h_psi_800_columns_ms=ms_h)
h.set_qq_nc(dnc)
ms = timed(lambda: h.s_psi_nc(lda, n, m // 2, psi, spsi))
out(item="s_psi_nc (400 spinors, re-uses becp%nc)", ms=ms)
h.close()
