"""CPU execution of the CUDA kernels' phase functions (tests/emul): the index /
twiddle / stick-table logic of q-e_b200/csrc/vloc_kernels.cuh, run serially
with plain g++, must reproduce the oracle. This is test infrastructure only -
the product path never runs on the CPU."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from oracle import descriptor as D
from oracle import hpsi_ref as R

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def emul():
    so = os.path.join(HERE, "emul", "libemul.so")
    subprocess.check_call(["sh", os.path.join(HERE, "emul", "build.sh")])
    L = ctypes.CDLL(so)
    P, I = ctypes.c_void_p, ctypes.c_int
    L.emul_fft1d.argtypes = [I, I, I, P, P, I, I]
    L.emul_vloc.argtypes = [I, I, I, I, P, P, P, I, I, I, P, P, P, P, I, I]
    L.emul_fullfft.argtypes = [I, I, I, I, I, P, I]
    L.emul_set_fast.argtypes = [I]
    return L


def p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


@pytest.mark.parametrize("n", [2, 3, 4, 5, 6, 8, 9, 10, 12, 15, 16, 18, 20, 24, 25, 27, 30, 32, 36, 45, 48, 50, 64, 72,
                               75, 80, 90, 96, 100, 108, 120, 125, 128, 135, 144, 180, 200, 216, 243, 256, 288, 300, 375])
def test_radix_passes_all_lengths(emul, n):
    rng = np.random.default_rng(n)
    nl = 5
    x = rng.standard_normal((nl, n)) + 1j * rng.standard_normal((nl, n))
    for sign, ref in ((1, np.fft.ifft(x, axis=1) * n), (-1, np.fft.fft(x, axis=1))):
        y = np.zeros_like(x)
        assert emul.emul_fft1d(n, sign, nl, p(x), p(y), 11, 10) == 0
        assert np.abs(y - ref).max() / np.abs(ref).max() < 1e-14


@pytest.mark.parametrize("maxr", [2, 3, 4, 5])
def test_radix_passes_small_radices_only(emul, maxr):
    """forces multi-pass plans out of the primitive butterflies (2,3,4,5)"""
    rng = np.random.default_rng(maxr)
    for n in (16, 27, 60, 125, 64, 81):
        if maxr == 2 and n % 2:
            continue
        if maxr == 2 and n in (60,):
            continue
        if maxr == 3 and n % 5 == 0:
            continue
        if maxr == 4 and n % 5 == 0:
            continue
        x = rng.standard_normal((3, n)) + 1j * rng.standard_normal((3, n))
        y = np.zeros_like(x)
        assert emul.emul_fft1d(n, 1, 3, p(x), p(y), 7, maxr) == 0
        ref = np.fft.ifft(x, axis=1) * n
        assert np.abs(y - ref).max() / np.abs(ref).max() < 1e-14


@pytest.mark.parametrize("name", ["si2_small", "ortho_small", "si2_small_gamma", "ortho_small_gamma", "si2", "al4"])
@pytest.mark.parametrize("m,batch", [(1, 1), (3, 2), (4, 3)])
@pytest.mark.parametrize("fast", [0, 1])
def test_vloc_phases_match_oracle(emul, name, m, batch, fast):
    """fast=1: the compile-time-plan, warp-mapped phase functions (vloc_fast.cuh) where a static plan exists"""
    emul.emul_set_fast(fast)
    d = D.config_descriptor(name)
    inp = R.synthetic_inputs(d, m, [4, 4], seed=m)
    lda, n = inp["lda"], inp["n"]
    rng = np.random.default_rng(0)
    nl_ = np.ascontiguousarray(d.nl, dtype=np.int32)
    nlm = np.ascontiguousarray(d.nlm, dtype=np.int32) if d.gamma_only else None
    igk = None if d.gamma_only else np.ascontiguousarray(d.igk, dtype=np.int32)
    for kin in (False, True):
        h0 = np.zeros((m, lda), complex) if kin else rng.standard_normal((m, lda)) + 0j
        ref = h0.copy()
        if kin:
            ref[:, :n] = d.g2kin[None, :n] * inp["psi"][:, :n]
        (R.vloc_psi_gamma if d.gamma_only else R.vloc_psi_k)(d, lda, n, m, inp["psi"], inp["vrs"], ref)
        out = h0.copy()
        g2 = np.ascontiguousarray(d.g2kin) if kin else None
        rc = emul.emul_vloc(d.nr1, d.nr2, d.nr3, n, p(nl_), p(nlm), p(igk), d.ngm, lda, m, p(inp["psi"]),
                            p(inp["vrs"]), p(g2), p(out), 96 if fast else 37, batch)
        assert rc == 0
        assert np.abs(out[:, :n] - ref[:, :n]).max() / np.abs(ref[:, :n]).max() < 1e-13


@pytest.mark.parametrize("shape", [(15, 15, 15), (24, 20, 18), (25, 27, 16), (45, 8, 50)])
def test_fullbox_fft_phases(emul, shape):
    nr1, nr2, nr3 = shape
    rng = np.random.default_rng(2)
    f = rng.standard_normal((2, nr3, nr2, nr1)) + 1j * rng.standard_normal((2, nr3, nr2, nr1))
    g = f.copy()
    assert emul.emul_fullfft(nr1, nr2, nr3, 2, 1, p(g), 19) == 0
    ref = np.fft.ifftn(f, axes=(1, 2, 3)) * (nr1 * nr2 * nr3)
    assert np.abs(g - ref).max() / np.abs(ref).max() < 1e-13
    assert emul.emul_fullfft(nr1, nr2, nr3, 2, -1, p(g), 19) == 0
    assert np.abs(g - f).max() / np.abs(f).max() < 1e-13


@pytest.mark.parametrize("name,m", [("ausurf112", 1), ("h2o64", 2), ("si512", 1)])
def test_vloc_fast_phases_full_baseline_grids(emul, name, m):
    """BASELINE grids (125x64x200 k, 135^3 and 144^3 Gamma) through the compile-time-plan phase functions."""
    emul.emul_set_fast(1)
    d = D.config_descriptor(name)
    inp = R.synthetic_inputs(d, m, [4, 4], seed=1)
    lda, n = inp["lda"], inp["n"]
    ref = np.zeros((m, lda), complex)
    (R.vloc_psi_gamma if d.gamma_only else R.vloc_psi_k)(d, lda, n, m, inp["psi"], inp["vrs"], ref)
    out = np.zeros((m, lda), complex)
    nl_ = np.ascontiguousarray(d.nl, dtype=np.int32)
    nlm = np.ascontiguousarray(d.nlm, dtype=np.int32) if d.gamma_only else None
    igk = None if d.gamma_only else np.ascontiguousarray(d.igk, dtype=np.int32)
    rc = emul.emul_vloc(d.nr1, d.nr2, d.nr3, n, p(nl_), p(nlm), p(igk), d.ngm, lda, m, p(inp["psi"]), p(inp["vrs"]),
                        None, p(out), 160, 2)
    assert rc == 0
    assert np.abs(out[:, :n] - ref[:, :n]).max() / np.abs(ref[:, :n]).max() < 1e-13
    emul.emul_set_fast(0)


def _run_plane_emul(emul, name, m, nthr, batch, mode=2):
    emul.emul_set_fast(mode)
    try:
        d = D.config_descriptor(name)
        inp = R.synthetic_inputs(d, m, [4, 4], seed=2)
        lda, n = inp["lda"], inp["n"]
        ref = np.zeros((m, lda), complex)
        (R.vloc_psi_gamma if d.gamma_only else R.vloc_psi_k)(d, lda, n, m, inp["psi"], inp["vrs"], ref)
        out = np.zeros((m, lda), complex)
        nl_ = np.ascontiguousarray(d.nl, dtype=np.int32)
        nlm = np.ascontiguousarray(d.nlm, dtype=np.int32) if d.gamma_only else None
        igk = None if d.gamma_only else np.ascontiguousarray(d.igk, dtype=np.int32)
        rc = emul.emul_vloc(d.nr1, d.nr2, d.nr3, n, p(nl_), p(nlm), p(igk), d.ngm, lda, m, p(inp["psi"]),
                            p(inp["vrs"]), None, p(out), nthr, batch)
        assert rc == 0
        assert np.abs(out[:, :n] - ref[:, :n]).max() / np.abs(ref[:, :n]).max() < 1e-13
    finally:
        emul.emul_set_fast(0)


@pytest.mark.parametrize("name", ["si2_small", "si2_small_gamma", "si2", "al4"])
@pytest.mark.parametrize("nthr", [32, 96, 512])
def test_plane_phases_match_oracle(emul, name, nthr):
    """the fused plane pass (vloc_plane.cuh: y | x*V*x | y on one shared-memory plane) on the small grids with an
    instantiated (x,y) plan pair, with fewer, as many and more warps than line groups"""
    _run_plane_emul(emul, name, 3, nthr, 2)


@pytest.mark.parametrize("name", ["si2_small", "si2_small_gamma", "si2", "al4"])
@pytest.mark.parametrize("nthr", [32, 96, 512])
def test_plane2_phases_match_oracle(emul, name, nthr):
    """the second form of the plane pass (vloc_plane2.cuh: compile-time phases, flat task list dealt to the warps,
    column masks) on the small grids"""
    _run_plane_emul(emul, name, 3, nthr, 2, mode=3)


@pytest.mark.parametrize("nthr", [512, 352, 64])
def test_plane2_phases_ausurf_grid(emul, nthr):
    """BASELINE AUSURF112 grid through the plane2 pass with the default plans (125 = 25*5, 64 = 8*8); 16, 11, 2 warps"""
    _run_plane_emul(emul, "ausurf112", 1, nthr, 1, mode=3)


@pytest.mark.parametrize("nthr", [800, 1024, 96])
def test_plane2_phases_ausurf_grid_555(emul, monkeypatch, nthr):
    """AUSURF112 grid through the plane2 pass with the low-register x plan 125 = 5*5*5 (the kernel's 25 warps; 32; 3)"""
    monkeypatch.setenv("QEB_XPLAN", "5,5,5")
    _run_plane_emul(emul, "ausurf112", 1, nthr, 1, mode=3)


def test_plane_phases_ausurf_grid(emul):
    """BASELINE AUSURF112 grid (125 x 64 plane, 200 planes) through the plane pass with the kernel's 16 warps"""
    _run_plane_emul(emul, "ausurf112", 1, 512, 1)


def test_plane_and_z_low_register_plans_ausurf(emul, monkeypatch):
    """QEB_XPLAN/QEB_YPLAN/QEB_ZPLAN overrides: 125 = 5x5x5, 64 = 4x4x4, 200 = 8x5x5 (the low-register static
    plans of fft_static.cuh) through the plane pass and the fast z passes on the AUSURF112 grid"""
    monkeypatch.setenv("QEB_XPLAN", "5,5,5")
    monkeypatch.setenv("QEB_YPLAN", "4,4,4")
    monkeypatch.setenv("QEB_ZPLAN", "8,5,5")
    _run_plane_emul(emul, "ausurf112", 1, 1024, 1)


@pytest.mark.parametrize("name", ["si2_small", "si2_small_gamma", "si2", "ortho_small"])
@pytest.mark.parametrize("fast", [0, 1])
def test_sum_band_density_phases(emul, name, fast):
    """the density-accumulation phases of qeb_sum_band_rho (vloc_kernels.cuh xacc_*, vloc_fast.cuh xacc_fast_phase) on
    the CPU against the oracle restatement of the sum_band loop (PW/src/sum_band.f90:326-647)"""
    emul.emul_sumband.argtypes = [ctypes.c_int] * 4 + [ctypes.c_void_p] * 3 + [ctypes.c_int] * 3 + [ctypes.c_void_p] * 3 + [ctypes.c_int] * 2
    emul.emul_set_fast(fast)
    try:
        d = D.config_descriptor(name)
        m = 5
        inp = R.synthetic_inputs(d, m, [4, 4], seed=3)
        lda, n = inp["lda"], inp["n"]
        rng = np.random.default_rng(1)
        w = np.ascontiguousarray(rng.uniform(0.0, 2.0, m))
        rho0 = rng.uniform(0.0, 1.0, d.nnr)
        ref = R.sum_band_rho(d, lda, n, m, inp["psi"], w, rho0.copy())
        rho = rho0.copy()
        nl_ = np.ascontiguousarray(d.nl, dtype=np.int32)
        nlm = np.ascontiguousarray(d.nlm, dtype=np.int32) if d.gamma_only else None
        igk = None if d.gamma_only else np.ascontiguousarray(d.igk, dtype=np.int32)
        rc = emul.emul_sumband(d.nr1, d.nr2, d.nr3, n, p(nl_), p(nlm), p(igk), d.ngm, lda, m, p(inp["psi"]), p(w), p(rho),
                               96 if fast else 37, 2)
        assert rc == 0
        assert np.abs(rho - ref).max() / np.abs(ref).max() < 1e-13
    finally:
        emul.emul_set_fast(0)


@pytest.mark.parametrize("name,fast,m", [("si2_small", 0, 3), ("si2", 0, 3), ("ortho_small", 0, 3), ("si2_small", 1, 3),
                                         ("si2", 1, 3), ("ortho_small", 1, 3), ("ausurf112", 1, 1)])
def test_vloc_nc_phases(emul, name, fast, m):
    """the spinor x pass (vloc_kernels.cuh xspin_phase: x DIF | 2x2 mix | x DIT on both components) inside the pass
    pipeline, on the CPU, against the oracle restatement of vloc_psi_nc (PW/src/vloc_psi_acc.f90:269-369, domag)"""
    emul.emul_vloc_nc.argtypes = [ctypes.c_int] * 4 + [ctypes.c_void_p] * 2 + [ctypes.c_int] * 3 + [ctypes.c_void_p] * 3 + [ctypes.c_int] * 2
    emul.emul_set_fast(fast)
    try:
        d = D.config_descriptor(name)
        inp = R.synthetic_inputs(d, 2 * m, [4, 4], seed=5)
        lda, n = inp["lda"], inp["n"]
        rng = np.random.default_rng(11)
        v4 = np.ascontiguousarray(rng.uniform(-1.0, 1.0, (4, d.nnr)))
        psi = np.ascontiguousarray(inp["psi"]).reshape(m, 2 * lda)
        h0 = (rng.standard_normal((m, 2 * lda)) + 1j * rng.standard_normal((m, 2 * lda)))
        ref = R.vloc_psi_nc(d, lda, n, m, psi, v4, h0.copy())
        out = np.ascontiguousarray(h0.copy())
        nl_ = np.ascontiguousarray(d.nl, dtype=np.int32)
        igk = np.ascontiguousarray(d.igk, dtype=np.int32)
        rc = emul.emul_vloc_nc(d.nr1, d.nr2, d.nr3, n, p(nl_), p(igk), d.ngm, lda, m, p(psi), p(v4), p(out),
                               (160 if d.nr1 > 30 else 96) if fast else 37, 2)
        assert rc == 0
        for b in range(m):
            for ipol in range(2):
                sl = slice(ipol * lda, ipol * lda + n)
                assert np.abs(out[b, sl] - ref[b, sl]).max() / np.abs(ref[b, sl]).max() < 1e-13
                pad = slice(ipol * lda + n, (ipol + 1) * lda)
                assert np.array_equal(out[b, pad], h0[b, pad])
    finally:
        emul.emul_set_fast(0)
