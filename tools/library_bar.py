#!/usr/bin/env python3
"""The same-box GPU LIBRARY bar: the reference's own GPU composition of h_psi - cuFFT
batched Z2Z over the FULL box + cuBLAS ZGEMM - timed on this B200 with the bench's inputs. Benchmark infrastructure
only: nothing under q-e_b200/ imports it, it uses torch (torch.fft = cuFFT, torch.matmul = cuBLAS) and no kernel of
libqeb200.

Composition restated (file:line relative to the QE 7.5 tree):
  kinetic        hpsi = g2kin * psi                                   PW/src/h_psi_gpu.f90:138-147
  per group of many_fft = 16 bands (Modules/control_flags.f90:285)    PW/src/vloc_psi_acc.f90:197-229
    zero-fill + scatter   psic = 0; psic(nl(igk)) = psi               FFTXlib/src/fft_helper_subroutines.f90:476-485
    invfft 'Wave'         full-box batched Z2Z (cfft3d_gpu: "outperforms cfft3ds_gpu")
                                                                      FFTXlib/src/fft_scalar.cuFFT.f90:523-654,695-698
    psic = psic * v(r)                                                PW/src/vloc_psi_acc.f90:212-217
    fwfft 'Wave'          full-box batched Z2Z, scaled 1/N
    gather + accumulate   hpsi += psic(nl(igk))                       PW/src/vloc_psi_acc.f90:222-227
  calbec         ZGEMM 'C','N' nkb x m x npw (MYZGEMM)                Modules/becmod.f90:389-390
  add_vuspsi     one nh x m x nh ZGEMM per atom, then ZGEMM 'N','N'   PW/src/add_vuspsi_acc.f90:178-200
Gamma: two bands per box (c1 + i c2 at nl, conj(c1 - i c2) at nlm), DGEMMs on the real views (becmod.f90:175-257,
add_vuspsi_acc.f90:90-130).
"""
from __future__ import annotations

import numpy as np

MANY_FFT = 16


class _HostEvent:  # CPU stand-in so that the composition itself can be checked against the oracle without a GPU
    def record(self):
        import time

        self.t = time.perf_counter()

    def elapsed_time(self, other):
        return (other.t - self.t) * 1e3


def _ev(torch):
    return torch.cuda.Event(enable_timing=True) if torch.cuda.is_available() else _HostEvent()


def _sync(torch):
    if torch.cuda.is_available():
        torch.cuda.synchronize()


def fp64_ceiling(torch, dev, n: int = 8192, reps: int = 3):
    """cuBLAS DGEMM and ZGEMM n^3 on this device: the empirical FP64 ceiling BASELINE.md section 2 asks for."""
    out = {}
    for name, dt, flops in (("dgemm", torch.float64, 2.0 * n ** 3), ("zgemm", torch.complex128, 8.0 * n ** 3)):
        a = torch.randn((n, n), device=dev, dtype=dt)
        b = torch.randn((n, n), device=dev, dtype=dt)
        c = torch.matmul(a, b)
        _sync(torch)
        e0, e1 = _ev(torch), _ev(torch)
        e0.record()
        for _ in range(reps):
            torch.matmul(a, b, out=c)
        e1.record()
        _sync(torch)
        out[f"{name}_{n}_tflops"] = flops * reps / (e0.elapsed_time(e1) * 1e-3) / 1e12
        del a, b, c
    return out


def measure(torch, d, psi, vkb, vrs, deeq, nh, ofs, reps: int = 2):
    """psi (m, lda, 2) and vkb (nkb, lda, 2) float64 CUDA tensors (the bench's), vrs/deeq numpy. Returns the bar:
    band-applications/s of the library composition and its per-section times."""
    dev = psi.device
    m, lda = psi.shape[0], psi.shape[1]
    n = d.npw
    nnr = d.nr1 * d.nr2 * d.nr3
    P = torch.view_as_complex(psi)
    V = torch.view_as_complex(vkb)
    nkb = V.shape[0]
    g2 = torch.from_numpy(np.ascontiguousarray(d.g2kin[:n])).to(dev)
    v = torch.from_numpy(np.ascontiguousarray(vrs)).to(dev)
    nl = np.asarray(d.nl, dtype=np.int64) - 1
    igk = (np.asarray(d.igk, dtype=np.int64) - 1) if not d.gamma_only else np.arange(n)
    idx = torch.from_numpy(nl[igk[:n]]).to(dev)
    idxm = torch.from_numpy(np.asarray(d.nlm, dtype=np.int64)[:n] - 1).to(dev) if d.gamma_only else None
    nhs = [int(x) for x in nh]
    D = [torch.from_numpy(np.ascontiguousarray(deeq[a, :nhs[a], :nhs[a]].T)).to(dev) for a in range(len(nhs))]
    if not d.gamma_only:
        D = [x.to(torch.complex128) for x in D]  # "deeq is real: copy it into a complex variable" (add_vuspsi.f90:148)
    H = torch.zeros((m, lda), dtype=torch.complex128, device=dev)
    box = (d.nr3, d.nr2, d.nr1)
    sec = {"kinetic": 0.0, "scatter": 0.0, "invfft": 0.0, "vmul": 0.0, "fwfft": 0.0, "gather": 0.0, "calbec": 0.0,
           "deeq": 0.0, "addvus": 0.0}

    def timed(key, fn):
        e0, e1 = _ev(torch), _ev(torch)
        e0.record()
        r = fn()
        e1.record()
        timed.pending.append((key, e0, e1))
        return r

    timed.pending = []

    def one_pass(acc):
        timed.pending = []
        def kin():
            H.zero_()
            H[:, :n] = g2[None, :] * P[:, :n]
        timed("kinetic", kin)
        step = 2 * MANY_FFT if d.gamma_only else MANY_FFT
        for b0 in range(0, m, step):
            nb = min(step, m - b0)
            if d.gamma_only:
                nbox = (nb + 1) // 2
                c1 = P[b0:b0 + nb:2, :n]
                c2 = P[b0 + 1:b0 + nb:2, :n]
                if c2.shape[0] < nbox:  # odd remainder: a single band in the last box
                    c2 = torch.cat([c2, torch.zeros_like(c1[:1])], 0)

                def sc():
                    psic = torch.zeros((nbox, nnr), dtype=torch.complex128, device=dev)
                    psic[:, idx] = c1 + 1j * c2
                    psic[:, idxm] = torch.conj(c1 - 1j * c2)
                    return psic
            else:
                nbox = nb

                def sc():
                    psic = torch.zeros((nbox, nnr), dtype=torch.complex128, device=dev)
                    psic[:, idx] = P[b0:b0 + nb, :n]
                    return psic
            psic = timed("scatter", sc)
            r = timed("invfft", lambda: torch.fft.ifftn(psic.view(nbox, *box), dim=(1, 2, 3), norm="forward"))
            timed("vmul", lambda: r.view(nbox, nnr).mul_(v[None, :]))
            g = timed("fwfft", lambda: torch.fft.fftn(r, dim=(1, 2, 3), norm="forward").view(nbox, nnr))
            if d.gamma_only:
                def ga():
                    a_, b_ = g[:, idx], g[:, idxm]
                    fp, fm = 0.5 * (a_ + b_), 0.5 * (a_ - b_)
                    H[b0:b0 + nb:2, :n] += torch.complex(fp.real, fm.imag)
                    k2 = nb // 2
                    if k2:
                        H[b0 + 1:b0 + nb:2, :n] += torch.complex(fp.imag[:k2], -fm.real[:k2])
                timed("gather", ga)
            else:
                def ga():
                    H[b0:b0 + nb, :n] += g[:, idx]
                timed("gather", ga)
        if nkb:
            if d.gamma_only:
                Pr = torch.view_as_real(P).view(m, 2 * lda)[:, :2 * n]  # the 2*npwx real view, no copy
                Vr = torch.view_as_real(V).view(nkb, 2 * lda)[:, :2 * n]

                def cb():
                    b = 2.0 * torch.matmul(Pr, Vr.T)
                    if d.gstart == 2:
                        b -= torch.outer(Pr[:, 0], Vr[:, 0])  # DGER at G = 0
                    return b
                becp = timed("calbec", cb)
            else:
                becp = timed("calbec", lambda: torch.matmul(P[:, :n], V[:, :n].conj().T))  # ZGEMM 'C','N'
            ps = torch.empty_like(becp)

            def dq():
                for a in range(len(nhs)):  # one small GEMM per atom (add_vuspsi_acc.f90:186-188)
                    o = int(ofs[a])
                    ps[:, o:o + nhs[a]] = torch.matmul(becp[:, o:o + nhs[a]], D[a].T)
            timed("deeq", dq)
            if d.gamma_only:
                Hr = torch.view_as_real(H).view(m, 2 * lda)[:, :2 * n]
                timed("addvus", lambda: Hr.addmm_(ps, Vr))  # DGEMM, beta = 1
            else:
                timed("addvus", lambda: H[:, :n].addmm_(ps, V[:, :n]))
        _sync(torch)
        if acc:
            for key, e0, e1 in timed.pending:
                sec[key] += e0.elapsed_time(e1)

    one_pass(False)  # warm-up: cuFFT plans, cuBLAS workspaces
    e0, e1 = _ev(torch), _ev(torch)
    e0.record()
    for _ in range(reps):
        one_pass(True)
    e1.record()
    _sync(torch)
    ms = e0.elapsed_time(e1) / reps
    sec = {k: v_ / reps for k, v_ in sec.items()}
    vloc_ms = sum(sec[k] for k in ("scatter", "invfft", "vmul", "fwfft", "gather"))
    gemm_ms = sec["calbec"] + sec["deeq"] + sec["addvus"]
    return {"value": m / (ms * 1e-3), "unit": "band-applications/s", "ms_per_step": ms, "sections_ms": sec,
            "vloc_us_per_band": 1e3 * vloc_ms / m, "projectors_us_per_band": 1e3 * gemm_ms / m,
            "many_fft": MANY_FFT, "hpsi": H,
            "composition": "cuFFT batched Z2Z full box (torch.fft, howmany 16) + zero-fill/scatter/V/gather elementwise "
                           "kernels + cuBLAS ZGEMM C,N + one ZGEMM per atom + ZGEMM N,N (torch.matmul): "
                           "fft_scalar.cuFFT.f90:523-654, vloc_psi_acc.f90:197-229, becmod.f90:389-390, "
                           "add_vuspsi_acc.f90:178-200"}


if __name__ == "__main__":
    import json
    import os
    import sys

    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import torch

    dev = torch.device("cuda", 0)
    print(json.dumps(fp64_ceiling(torch, dev)))
