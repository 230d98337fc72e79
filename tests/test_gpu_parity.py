"""GPU parity tests (run with -m gpu on a B200): the CUDA path, called through
the C ABI (ctypes mirror qe_b200.HPsi), against the CPU oracle on identical
seeded inputs. Tolerance: 1e-11 relative (north_star), FP64."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-11


@pytest.fixture(scope="module")
def mods():
    import torch

    if not torch.cuda.is_available():
        pytest.fail("GPU tests need a CUDA device; there is no CPU fallback")
    import qe_b200
    from oracle import descriptor as D
    from oracle import hpsi_ref as R

    return qe_b200, D, R, torch


def relerr(a, b):
    """Per-column error: max over the columns of the Fortran array (leading numpy index: one band, one projector
    row of becp, one FFT box) of max|a - b| / max|b| of THAT column, so that a column which is wrong relative to its
    own scale cannot hide behind a larger one (1-D arrays: one column)."""
    a = np.asarray(a)
    b = np.asarray(b)
    if a.ndim < 2:
        a, b = a.reshape(1, -1), b.reshape(1, -1)
    a2, b2 = a.reshape(a.shape[0], -1), b.reshape(b.shape[0], -1)
    num = np.abs(a2 - b2).max(axis=1)
    den = np.abs(b2).max(axis=1)
    with np.errstate(divide="ignore", invalid="ignore"):
        r = np.where(den > 0, num / den, np.where(num > 0, np.inf, 0.0))
    return float(r.max()) if r.size else 0.0


@pytest.mark.parametrize("shape", [(15, 15, 15), (25, 25, 25), (24, 20, 18), (27, 27, 27), (64, 45, 50), (75, 120, 180)])
def test_fft_roundtrip_and_numpy(mods, shape):
    qe, D, R, torch = mods
    nr1, nr2, nr3 = shape
    h = qe.HPsi()
    h.set_grid(nr1, nr2, nr3, nl=np.array([1], dtype=np.int32))
    rng = np.random.default_rng(1)
    how = 2
    f = rng.standard_normal((how, nr3, nr2, nr1)) + 1j * rng.standard_normal((how, nr3, nr2, nr1))
    g = f.copy()
    h.invfft("Rho", g, how)
    ref = np.fft.ifftn(f, axes=(1, 2, 3)) * (nr1 * nr2 * nr3)
    assert relerr(g, ref) < 1e-13
    h.fwfft("Rho", g, how)
    assert relerr(g, f) < 1e-13
    # device pointer path
    t = torch.from_numpy(f.copy()).cuda()
    h.fwfft("Wave", t, how)
    torch.cuda.synchronize()
    assert relerr(t.cpu().numpy(), np.fft.fftn(f, axes=(1, 2, 3)) / (nr1 * nr2 * nr3)) < 1e-13
    h.close()


def test_fft_known_answer_fft_test0(mods, golden_dir):
    """FFTXlib/examples/reference/fft_test0.out: analytic 4-coefficient transform on 75x120x180."""
    qe, D, R, torch = mods
    kat = json.load(open(os.path.join(golden_dir, "fft_test0.json")))
    nr = kat["grid"]
    box = np.zeros(nr[0] * nr[1] * nr[2], dtype=np.complex128)
    for key, (re_, im_) in kat["coeff"].items():
        i, j, k = (int(t) - 1 for t in key.split(","))
        box[i + j * nr[0] + k * nr[0] * nr[1]] = re_ + 1j * im_
    h = qe.HPsi()
    h.set_grid(nr[0], nr[1], nr[2], nl=np.array([1], dtype=np.int32))
    f = box.copy()
    h.invfft("Wave", f)
    b3 = f.reshape(nr[2], nr[1], nr[0])
    gold = np.array(kat["real_space_5x5x5_kji"]).reshape(5, 5, 5, 2)
    got = b3[:5, :5, :5]
    assert np.abs(got.real - gold[..., 0]).max() < 1.01e-6  # printed to 6 decimals
    assert np.abs(got.imag - gold[..., 1]).max() < 1.01e-6
    h.fwfft("Wave", f)
    assert np.abs(f - box).max() < 1e-14
    h.close()


SMALL = ["si2_small", "ortho_small", "si2_small_gamma", "ortho_small_gamma"]


@pytest.mark.parametrize("name", SMALL)
@pytest.mark.parametrize("m", [1, 2, 5, 8])
def test_vloc_psi(mods, name, m):
    qe, D, R, torch = mods
    d = D.config_descriptor(name)
    inp = R.synthetic_inputs(d, m, [4, 4], seed=m)
    lda, n = inp["lda"], inp["n"]
    rng = np.random.default_rng(5)
    h0 = rng.standard_normal((m, lda)) + 1j * rng.standard_normal((m, lda))
    ref = h0.copy()
    (R.vloc_psi_gamma if d.gamma_only else R.vloc_psi_k)(d, lda, n, m, inp["psi"], inp["vrs"], ref)
    h = qe.HPsi(batch=3)
    h.setup_from(d, inp)
    out = h0.copy()
    (h.vloc_psi_gamma if d.gamma_only else h.vloc_psi_k)(lda, n, m, inp["psi"], inp["vrs"], out)
    assert relerr(out[:, :n], ref[:, :n]) < TOL
    assert np.array_equal(out[:, n:], h0[:, n:])  # rows beyond n untouched
    # device pointers, potential from the context
    psi_t = torch.from_numpy(np.nan_to_num(inp["psi"])).cuda()
    out_t = torch.from_numpy(h0.copy()).cuda()
    (h.vloc_psi_gamma if d.gamma_only else h.vloc_psi_k)(lda, n, m, psi_t, None, out_t)
    torch.cuda.synchronize()
    assert relerr(out_t.cpu().numpy()[:, :n], ref[:, :n]) < TOL
    h.close()


@pytest.mark.parametrize("name", ["si2_small", "ortho_small", "si2", "al4"])
@pytest.mark.parametrize("m", [1, 2, 5])
def test_vloc_psi_nc(mods, name, m):
    """noncollinear spinors through the 2x2 potential (vloc_psi_nc, domag): qeb_vloc_psi_nc against the oracle
    restatement of PW/src/vloc_psi_acc.f90:269-369"""
    qe, D, R, torch = mods
    d = D.config_descriptor(name)
    inp = R.synthetic_inputs(d, 2 * m, [4, 4], seed=m)
    lda, n = inp["lda"], inp["n"]
    rng = np.random.default_rng(17)
    v4 = np.ascontiguousarray(rng.uniform(-1.0, 1.0, (4, d.nnr)))
    psi = np.ascontiguousarray(inp["psi"]).reshape(m, 2 * lda)
    h0 = rng.standard_normal((m, 2 * lda)) + 1j * rng.standard_normal((m, 2 * lda))
    ref = R.vloc_psi_nc(d, lda, n, m, psi, v4, h0.copy())
    h = qe.HPsi(batch=3)  # odd batch: one spinor per launch
    h.setup_from(d, inp)
    out = h0.copy()
    h.vloc_psi_nc(lda, n, m, psi, v4, out)
    live = np.zeros(2 * lda, bool)
    live[:n] = live[lda:lda + n] = True
    assert relerr(out[:, live], ref[:, live]) < TOL
    assert np.array_equal(out[:, ~live], h0[:, ~live])  # rows beyond n of both components untouched
    # device pointers, potential from the context, larger batch
    h2 = qe.HPsi(batch=8)
    h2.setup_from(d, inp)
    h2.set_vloc_nc(v4)
    psi_t = torch.from_numpy(np.nan_to_num(psi)).cuda()
    out_t = torch.from_numpy(h0.copy()).cuda()
    h2.vloc_psi_nc(lda, n, m, psi_t, None, out_t)
    torch.cuda.synchronize()
    assert relerr(out_t.cpu().numpy()[:, live], ref[:, live]) < TOL
    # v2 = v3 = v4 = 0 reduces to the scalar path on 2m columns (the non-magnetic branch, :342-348)
    v1 = np.zeros_like(v4)
    v1[0] = inp["vrs"]
    a = h0.copy()
    h.vloc_psi_nc(lda, n, m, psi, v1, a)
    b = h0.copy().reshape(2 * m, lda)
    h.vloc_psi_k(lda, n, 2 * m, psi.reshape(2 * m, lda), inp["vrs"], b)
    assert relerr(a.reshape(2 * m, lda)[:, :n], b[:, :n]) < TOL
    h.close()
    h2.close()


@pytest.mark.parametrize("name", ["si2_small", "ortho_small", "si2", "al4"])
@pytest.mark.parametrize("m", [1, 3, 6])
def test_h_psi_nc(mods, name, m):
    """noncollinear h_psi with magnetisation (kinetic on both components, vloc_psi_nc, calbec_nc, add_vuspsi_nc with a
    Hermitian spin-dependent deeq_nc) against the oracle restatement; becp%nc left in the context"""
    qe, D, R, torch = mods
    d = D.config_descriptor(name)
    c = d.extra["config"]
    nh = c.get("nh_list") or [c["nh"]] * c["nat"]
    inp = R.synthetic_inputs(d, 2 * m, nh, seed=6)
    lda, n = inp["lda"], inp["n"]
    rng = np.random.default_rng(29)
    v4 = np.ascontiguousarray(rng.uniform(-1.0, 1.0, (4, d.nnr)))
    dnc = R.synthetic_deeq_nc(inp["nh_at"], seed=m)
    psi = np.ascontiguousarray(inp["psi"]).reshape(m, 2 * lda)
    ref, becp_ref = R.h_psi_nc(d, lda, n, m, np.nan_to_num(psi), v4, np.nan_to_num(inp["vkb"]), dnc, inp["nh_at"],
                               inp["ofsbeta"])
    h = qe.HPsi(batch=4)
    h.setup_from(d, inp)
    h.set_vloc_nc(v4)
    h.set_nonlocal_nc(dnc)
    out = np.full((m, 2 * lda), np.nan + 1j * np.nan)
    h.h_psi_nc(lda, n, m, psi, out)
    assert relerr(out, ref) < TOL  # includes the zeroed rows n+1..lda of both components
    becp = h.get_becp(2 * m).reshape(m, 2, -1)
    assert relerr(becp, becp_ref) < TOL
    # device pointers
    psi_t = torch.from_numpy(np.nan_to_num(psi)).cuda()
    out_t = torch.full((m, 2 * lda, 2), float("nan"), dtype=torch.float64, device="cuda")
    h.h_psi_nc(lda, n, m, psi_t, out_t)
    torch.cuda.synchronize()
    assert relerr(torch.view_as_complex(out_t).cpu().numpy(), ref) < TOL
    # s_psi_nc with a spin-orbit-like 2x2 overlap, from the becp%nc in the context
    qnc = R.synthetic_deeq_nc(inp["nh_at"], seed=100 + m)
    sref = R.s_psi_nc(lda, n, m, np.nan_to_num(psi), becp_ref, np.nan_to_num(inp["vkb"]), qnc, inp["nh_at"], inp["ofsbeta"])
    h.h_psi_nc(lda, n, m, psi, out)
    h.set_qq_nc(qnc)
    sp = np.zeros((m, 2 * lda), complex)
    h.s_psi_nc(lda, n, m, np.nan_to_num(psi), sp)
    assert relerr(sp, sref) < TOL
    # Hermitian spinor operator: <a|H b> = conj(<b|H a>)
    if m >= 2:
        live = np.zeros(2 * lda, bool)
        live[:n] = live[lda:lda + n] = True
        P = np.nan_to_num(psi)[:, live]
        A = P.conj() @ out[:, live].T
        assert np.abs(A - A.conj().T).max() / np.abs(A).max() < 1e-11
    h.close()


@pytest.mark.parametrize("generic", [0, 1])
def test_vloc_psi_nc_full_grid_ausurf(mods, generic, monkeypatch):
    """the spinor x pass at the AUSURF112 grid (125 = 25*5: the in-register mix between the last DIF and the first DIT
    butterflies of both components, k_vloc_xspin_f) and the runtime-plan kernel (QEB_GENERIC=1) against the oracle"""
    qe, D, R, torch = mods
    if generic:
        monkeypatch.setenv("QEB_GENERIC", "1")
    d = D.config_descriptor("ausurf112")
    m = 2
    inp = R.synthetic_inputs(d, 2 * m, [4], seed=4, lda=d.npw + 3)
    lda, n = inp["lda"], inp["n"]
    rng = np.random.default_rng(23)
    v4 = np.ascontiguousarray(rng.uniform(-1.0, 1.0, (4, d.nnr)))
    psi = np.ascontiguousarray(inp["psi"]).reshape(m, 2 * lda)
    h0 = np.zeros((m, 2 * lda), complex)
    ref = R.vloc_psi_nc(d, lda, n, m, psi, v4, h0.copy())
    h = qe.HPsi()
    h.setup_from(d, inp)
    out = h0.copy()
    h.vloc_psi_nc(lda, n, m, psi, v4, out)
    live = np.zeros(2 * lda, bool)
    live[:n] = live[lda:lda + n] = True
    assert relerr(out[:, live], ref[:, live]) < TOL
    h.close()


@pytest.mark.parametrize("name", SMALL + ["si2", "al4"])
@pytest.mark.parametrize("m", [1, 3, 4, 10])
def test_h_psi(mods, name, m):
    qe, D, R, torch = mods
    d = D.config_descriptor(name)
    c = d.extra["config"]
    nh = c.get("nh_list") or [c["nh"]] * c["nat"]
    inp = R.synthetic_inputs(d, m, nh, seed=3)
    lda, n = inp["lda"], inp["n"]
    ref, becp_ref = R.h_psi(d, lda, n, m, inp["psi"], inp["vrs"], inp["vkb"], inp["deeq"], inp["nh_at"], inp["ofsbeta"])
    h = qe.HPsi()
    h.setup_from(d, inp)
    out = np.full((m, lda), np.nan + 1j * np.nan)
    h.h_psi(lda, n, m, inp["psi"], out)  # psi rows n..lda are NaN-poisoned: must be ignored
    assert relerr(out, ref) < TOL
    assert np.all(out[:, n:] == 0)
    becp = h.get_becp(m)
    assert relerr(becp, becp_ref) < TOL
    if d.gamma_only and d.gstart == 2:
        assert np.all(out[:, 0].imag == 0)
    h.close()


@pytest.mark.parametrize("name", ["ortho_small", "ortho_small_gamma"])
def test_calbec_add_vuspsi(mods, name):
    qe, D, R, torch = mods
    d = D.config_descriptor(name)
    m = 6
    # a NON-symmetric deeq: pins the (ih,jh) orientation of deeq(ih,jh,na) against add_vuspsi.f90:151-155
    inp = R.synthetic_inputs(d, m, [8, 8, 8], seed=11, poison=False, symmetric=False)
    lda, n = inp["lda"], inp["n"]
    h = qe.HPsi()
    h.setup_from(d, inp)
    if d.gamma_only:
        ref = R.calbec_gamma(n, inp["vkb"], inp["psi"], m, d.gstart)
        out = np.zeros((m, inp["nkb"]))
    else:
        ref = R.calbec_k(n, inp["vkb"], inp["psi"], m)
        out = np.zeros((m, inp["nkb"]), dtype=complex)
    h.calbec(n, inp["vkb"], inp["psi"], out, m, lda=lda)  # explicit beta (host)
    assert relerr(out, ref) < TOL
    out2 = np.zeros_like(out)
    h.calbec(n, None, inp["psi"], out2, m, lda=lda)  # the context's vkb
    assert relerr(out2, ref) < TOL
    # m == 1 edge (the reference switches to GEMV, becmod.f90:468)
    o1 = np.zeros_like(out[:1])
    h.calbec(n, None, inp["psi"][:1], o1, 1, lda=lda)
    assert relerr(o1, ref[:1]) < TOL
    ps = R.apply_deeq(ref, inp["deeq"], inp["nh_at"], inp["ofsbeta"])
    rng = np.random.default_rng(2)
    h0 = rng.standard_normal((m, lda)) + 1j * rng.standard_normal((m, lda))
    href = h0.copy()
    (R.add_vuspsi_gamma if d.gamma_only else R.add_vuspsi_k)(n, m, inp["vkb"], ps, href)
    hout = h0.copy()
    h.add_vuspsi(lda, n, m, hout, ref.copy())
    assert relerr(hout, href) < TOL
    h.close()


def test_errors(mods):
    qe, D, R, torch = mods
    h = qe.HPsi()
    with pytest.raises(qe.QebError):
        h.set_grid(14, 15, 15, nl=np.array([1], dtype=np.int32))  # factor 7 not allowed (fft_support.f90)
    d = D.config_descriptor("si2_small")
    inp = R.synthetic_inputs(d, 2, [4, 4])
    with pytest.raises(qe.QebError):
        h.h_psi(inp["lda"], inp["n"], 2, inp["psi"], inp["psi"].copy())  # no grid/k-point yet
    h.setup_from(d, inp)
    with pytest.raises(qe.QebError):
        h.h_psi(inp["lda"], inp["n"] - 1, 2, inp["psi"], inp["psi"].copy())  # n != npw of the context
    with pytest.raises(qe.QebError):
        h.vloc_psi_gamma(inp["lda"], inp["n"], 2, inp["psi"], None, inp["psi"].copy())  # k-point context
    h.close()


def test_band_groups_match_full_call(mods):
    """h_psi over band groups (PW/src/h_psi.f90:53-67): the column blocks of divide_all, applied
    independently, reproduce the full call bit for bit (no collective on one GPU)."""
    qe, D, R, torch = mods
    d = D.config_descriptor("ortho_small")
    m = 7
    inp = R.synthetic_inputs(d, m, [8, 8, 8], seed=4, poison=False)
    lda, n = inp["lda"], inp["n"]
    h = qe.HPsi()
    h.setup_from(d, inp)
    full = np.zeros((m, lda), dtype=complex)
    h.h_psi(lda, n, m, inp["psi"], full)
    parts = np.zeros_like(full)
    for g in range(3):
        h.h_psi_bgrp(lda, n, m, inp["psi"], parts, 3, g)
    assert relerr(parts, full) < 1e-14
    h.close()


@pytest.mark.parametrize("name,m,nat_used", [("h2o64", 3, 6), ("h2o64", 8, 3), ("ausurf112", 2, 4), ("si512", 5, 6), ("si512", 16, 2)])
def test_h_psi_full_grid_reduced_projectors(mods, name, m, nat_used):
    """BASELINE grids at full size (135^3 Gamma, 125x64x200 k, 144^3 Gamma) against the oracle, with a reduced
    projector count so that the CPU oracle finishes in seconds; m = 5 at Gamma: two band pairs + the odd remainder."""
    qe, D, R, torch = mods
    d = D.config_descriptor(name)
    c = d.extra["config"]
    nh = (c.get("nh_list") or [c["nh"]] * c["nat"])[:nat_used]
    inp = R.synthetic_inputs(d, m, nh, seed=9, lda=d.npw + 5)
    lda, n = inp["lda"], inp["n"]
    ref, becp_ref = R.h_psi(d, lda, n, m, inp["psi"], inp["vrs"], inp["vkb"], inp["deeq"], inp["nh_at"], inp["ofsbeta"])
    h = qe.HPsi()
    h.setup_from(d, inp)
    out = np.full((m, lda), np.nan + 1j * np.nan)
    h.h_psi(lda, n, m, inp["psi"], out)
    assert relerr(out, ref) < TOL
    h.close()


def test_hermiticity_full_size_ausurf(mods):
    """Size-independent property at BASELINE size (AUSURF112 grid, all 2016 projectors, synthetic
    vkb generated on the device): <a|H b> = conj(<b|H a>) and linearity."""
    qe, D, R, torch = mods
    d = D.config_descriptor("ausurf112")
    n = d.npw
    lda = n
    nkb = 2016
    g = torch.Generator(device="cuda").manual_seed(0)
    vkb = torch.randn((nkb, lda, 2), generator=g, device="cuda", dtype=torch.float64) / np.sqrt(n)
    m = 8
    psi = torch.randn((m, lda, 2), generator=g, device="cuda", dtype=torch.float64)
    rng = np.random.default_rng(3)
    vrs = rng.uniform(-1, 0, d.nnr)
    nh = np.full(112, 18, dtype=np.int32)
    ofs = (np.arange(112) * 18).astype(np.int32)
    deeq = rng.uniform(-1, 1, (112, 18, 18))
    deeq = 0.5 * (deeq + deeq.transpose(0, 2, 1))
    h = qe.HPsi()
    h.set_grid(d.nr1, d.nr2, d.nr3, d.nl)
    h.set_kpoint(n, lda, d.igk, d.g2kin, vkb, nkb, d.gstart)
    h.set_vloc(vrs)
    h.set_nonlocal(nh, ofs, deeq)
    hpsi = torch.empty_like(psi)
    h.h_psi(lda, n, m, psi, hpsi)
    torch.cuda.synchronize()
    P = torch.view_as_complex(psi)
    Hm = torch.view_as_complex(hpsi)
    A = P.conj() @ Hm.T  # A[a,b] = <a|H b>
    herm = (A - A.conj().T).abs().max().item() / A.abs().max().item()
    assert herm < 1e-11
    # linearity: H(psi0 + 2 psi1) = H psi0 + 2 H psi1
    comb = (P[0] + 2.0 * P[1]).contiguous()
    cin = torch.view_as_real(comb).contiguous().reshape(1, lda, 2)
    cout = torch.empty_like(cin)
    h.h_psi(lda, n, 1, cin, cout)
    torch.cuda.synchronize()
    lin = (torch.view_as_complex(cout)[0] - (Hm[0] + 2.0 * Hm[1])).abs().max().item() / Hm.abs().max().item()
    assert lin < 1e-12
    h.close()


@pytest.mark.parametrize("name", ["ortho_small", "ortho_small_gamma"])
@pytest.mark.parametrize("m,chunk,sub", [(11, 4, 0), (12, 6, 0), (7, 2, 0), (12, 6, 2), (11, 8, 4), (9, 9, 2), (13, 4, 2)])
def test_h_psi_host_pipeline(mods, name, m, chunk, sub, monkeypatch):
    """Host psi/hpsi through the chunked three-stream pipeline (qeb_set_pipeline): same result as the oracle,
    pad rows zeroed, becp of all columns kept; pageable numpy arrays (page-locking is explicit, see test_pin_host_explicit).
    sub > 0: the first chunk is copied in sub-chunks with the V_loc passes following them (QEB_PIPE_SUB)."""
    qe, D, R, torch = mods
    if sub:
        monkeypatch.setenv("QEB_PIPE_SUB", str(sub))
        monkeypatch.setenv("QEB_PIPE_TAIL", str(sub))  # add_vuspsi + copy-out of the last chunks in sub-chunks
    d = D.config_descriptor(name)
    inp = R.synthetic_inputs(d, m, [8, 8, 8], seed=11)
    lda, n = inp["lda"], inp["n"]
    ref, becp_ref = R.h_psi(d, lda, n, m, inp["psi"], inp["vrs"], inp["vkb"], inp["deeq"], inp["nh_at"], inp["ofsbeta"])
    h = qe.HPsi()
    h.setup_from(d, inp)
    h.set_pipeline(chunk, 2)
    for _ in range(2):
        out = np.full((m, lda), np.nan + 1j * np.nan)
        h.h_psi(lda, n, m, inp["psi"], out)
        assert relerr(out, ref) < TOL
        assert np.all(out[:, n:] == 0)
        assert relerr(h.get_becp(m), becp_ref) < TOL
    h.close()


@pytest.mark.parametrize("name", SMALL + ["si2", "al4"])
@pytest.mark.parametrize("m", [1, 4, 9])
def test_s_psi_after_h_psi(mods, name, m):
    """s_psi (PW/src/s_psi.f90:10-262) right after h_psi, as cegterg calls them (cegterg.f90:192-196): it re-uses
    the becp that h_psi left behind; every lda row of psi is copied (threaded_memcpy, :118)."""
    qe, D, R, torch = mods
    d = D.config_descriptor(name)
    c = d.extra["config"]
    nh = c.get("nh_list") or [c["nh"]] * c["nat"]
    inp = R.synthetic_inputs(d, m, nh, seed=5, poison=False)
    lda, n = inp["lda"], inp["n"]
    qq = R.synthetic_qq(nh, seed=5, symmetric=(m != 4))  # m = 4: general qq, pins the (ih,jh) orientation
    _, becp_ref = R.h_psi(d, lda, n, m, inp["psi"], inp["vrs"], inp["vkb"], inp["deeq"], inp["nh_at"], inp["ofsbeta"])
    ref = R.s_psi(d, lda, n, m, inp["psi"], becp_ref, inp["vkb"], qq, inp["nh_at"], inp["ofsbeta"])
    h = qe.HPsi()
    h.setup_from(d, inp)
    out = np.empty((m, lda), dtype=complex)
    # norm-conserving run (no qq): plain copy
    h.h_psi(lda, n, m, inp["psi"], out)
    sp = np.full((m, lda), np.nan + 1j * np.nan)
    h.s_psi(lda, n, m, inp["psi"], sp)
    assert np.array_equal(sp, inp["psi"])
    # ultrasoft S
    h.set_qq(qq)
    h.s_psi(lda, n, m, inp["psi"], sp)
    assert relerr(sp, ref) < TOL
    assert np.array_equal(sp[:, n:], inp["psi"][:, n:])
    # device pointers
    psi_t = torch.from_numpy(inp["psi"]).cuda()
    sp_t = torch.empty_like(psi_t)
    hp_t = torch.empty_like(psi_t)
    h.h_psi(lda, n, m, psi_t, hp_t)
    h.s_psi(lda, n, m, psi_t, sp_t)
    torch.cuda.synchronize()
    assert relerr(sp_t.cpu().numpy(), ref) < TOL
    h.close()


def test_s_psi_needs_becp(mods):
    qe, D, R, torch = mods
    d = D.config_descriptor("si2_small")
    inp = R.synthetic_inputs(d, 3, [4, 4], poison=False)
    h = qe.HPsi()
    h.setup_from(d, inp)
    h.set_qq(R.synthetic_qq([4, 4]))
    with pytest.raises(qe.QebError):
        h.s_psi(inp["lda"], inp["n"], 3, inp["psi"], inp["psi"].copy())  # no h_psi / calbec before
    out = np.zeros((3, inp["nkb"]), dtype=complex)
    h.calbec(inp["n"], None, inp["psi"], out, 3, lda=inp["lda"])
    h.s_psi(inp["lda"], inp["n"], 3, inp["psi"], inp["psi"].copy())
    h.close()


@pytest.mark.parametrize("name,nsp,lmaxkb,nat", [("si2_small", 1, 1, 2), ("ortho_small", 2, 2, 3), ("si2", 2, 3, 2),
                                                 ("ortho_small_gamma", 1, 2, 2)])
def test_g2_kin_init_us_2_on_device(mods, name, nsp, lmaxkb, nat):
    """g2_kin (PW/src/g2_kin.f90) and init_us_2 (upflib/init_us_2_acc.f90, ylmr2, interp_beta) built on
    the device, then consumed by h_psi without any vkb / g2kin upload."""
    qe, D, R, torch = mods
    from oracle import uspp_ref as U

    d = D.config_descriptor(name)
    us = U.synthetic_uspp(nsp, lmaxkb, seed=nat)
    rng = np.random.default_rng(11)
    tau = rng.uniform(-0.5, 0.5, (nat, 3))
    ityp = np.array([(i % nsp) + 1 for i in range(nat)], dtype=np.int32)
    nr_dense = (d.nr1 + 4, d.nr2 + 2, d.nr3 + 6)  # the dense grid only sizes the eigts tables
    eig = U.eigts(d.bg, tau, nr_dense)
    tpiba = 2.0 * np.pi / d.alat
    n = d.npw
    npwx = n + 3
    igk = np.arange(1, n + 1, dtype=np.int32) if d.gamma_only else d.igk
    vkb_ref = U.init_us_2(n, npwx, igk, d.xk, tau, ityp, tpiba, nr_dense, eig, d.mill, d.g, us)
    nkb = vkb_ref.shape[0]
    g2_ref = U.g2_kin(d.g, igk, d.xk, d.tpiba2)
    h = qe.HPsi()
    h.set_grid(d.nr1, d.nr2, d.nr3, d.nl, d.nlm if d.gamma_only else None, d.gamma_only)
    h.set_kpoint(n, npwx, None if d.gamma_only else d.igk, None, None, nkb, d.gstart)
    h.set_gvectors(d.g, d.mill)
    g2 = np.zeros(n)
    h.g2_kin(d.xk, d.tpiba2, out=g2)
    assert relerr(g2, g2_ref) < 1e-15
    g2m = np.zeros(n)  # modified kinetic functional (qcutz > 0)
    h.g2_kin(d.xk, d.tpiba2, 20.0, 5.0, 0.7, out=g2m)
    assert relerr(g2m, U.g2_kin(d.g, igk, d.xk, d.tpiba2, 20.0, 5.0, 0.7)) < 1e-14
    h.g2_kin(d.xk, d.tpiba2)
    h.set_beta_tables(us)
    h.set_structure(ityp, tau, nr_dense, eig)
    vkb = np.full((nkb, npwx), np.nan + 0j)
    h.init_us_2(d.xk, tpiba, out=vkb)
    assert relerr(vkb, vkb_ref) < 1e-13
    assert np.all(vkb[:, n:] == 0)
    # h_psi consumes the device-built state
    m = 3
    nh_at = [int(us["nh"][t - 1]) for t in sorted(ityp)]  # reference column order: atoms grouped by type
    inp = R.synthetic_inputs(d, m, nh_at, seed=2, lda=npwx)
    if d.gamma_only:  # the Gamma trick needs real beta(G=0); the synthetic tables do not guarantee it
        h.close()
        return
    ref, _ = R.h_psi(d, npwx, n, m, inp["psi"], inp["vrs"], vkb_ref, inp["deeq"], inp["nh_at"], inp["ofsbeta"])
    h.set_vloc(inp["vrs"])
    h.set_nonlocal(inp["nh_at"], inp["ofsbeta"], inp["deeq"])
    out = np.full((m, npwx), np.nan + 0j)
    h.h_psi(npwx, n, m, inp["psi"], out)
    assert relerr(out, ref) < TOL
    h.close()


def test_h_psi_without_state_fails(mods):
    qe, D, R, torch = mods
    d = D.config_descriptor("si2_small")
    inp = R.synthetic_inputs(d, 2, [4, 4], poison=False)
    h = qe.HPsi()
    h.set_grid(d.nr1, d.nr2, d.nr3, d.nl)
    h.set_kpoint(d.npw, inp["lda"], d.igk, None, None, 8, d.gstart)  # neither g2kin nor vkb yet
    h.set_vloc(inp["vrs"])
    with pytest.raises(qe.QebError):
        h.h_psi(inp["lda"], d.npw, 2, inp["psi"], inp["psi"].copy())
    with pytest.raises(qe.QebError):
        h.g2_kin(d.xk, d.tpiba2)  # no G-vector list
    h.close()


@pytest.mark.parametrize("name", ["si2_small", "si2_small_gamma"])
def test_hs_psi_and_hs_1psi(mods, name):
    """hs_psi / hs_1psi (PW/src/hs_psi.f90:7, hs_1psi.f90:7): H and S of the same block in one call"""
    qe, D, R, torch = mods
    d = D.config_descriptor(name)
    m = 3
    inp = R.synthetic_inputs(d, m, [4, 4], seed=23, poison=False)
    lda, n = inp["lda"], inp["n"]
    qq = R.synthetic_qq([4, 4], seed=23)
    h = qe.HPsi()
    h.setup_from(d, inp)
    h.set_qq(qq)
    href, becp = R.h_psi(d, lda, n, m, inp["psi"], inp["vrs"], inp["vkb"], inp["deeq"], inp["nh_at"], inp["ofsbeta"])
    sref = R.s_psi(d, lda, n, m, inp["psi"], becp, inp["vkb"], qq, inp["nh_at"], inp["ofsbeta"])
    hp, sp = np.full_like(inp["psi"], np.nan), np.full_like(inp["psi"], np.nan)
    h.hs_psi(lda, n, m, inp["psi"], hp, sp)
    assert relerr(hp[:, :n], href[:, :n]) < TOL and relerr(sp[:, :n], sref[:, :n]) < TOL
    h1, s1 = np.full_like(inp["psi"][:1], np.nan), np.full_like(inp["psi"][:1], np.nan)
    h.hs_psi(lda, n, 1, inp["psi"][1:2].copy(), h1, s1)  # hs_1psi
    assert relerr(h1[:, :n], href[1:2, :n]) < TOL and relerr(s1[:, :n], sref[1:2, :n]) < TOL
    h.close()


@pytest.mark.parametrize("name", ["ortho_small", "ortho_small_gamma"])
def test_davidson_building_blocks(mods, name):
    """the device pieces of cegterg/regterg - reduced matrices psi^H hpsi (cegterg.f90:209, regterg.f90:203-205),
    basis update psi vc (:353-395), g_psi preconditioner (PW/src/g_psi.f90) and the correction norms (:421) - against
    numpy, composed into one Rayleigh-Ritz step on the oracle's H."""
    qe, D, R, torch = mods
    d = D.config_descriptor(name)
    nb, notcnv = 7, 3
    inp = R.synthetic_inputs(d, nb, [8, 8, 8], seed=21, poison=False)
    lda, n = inp["lda"], inp["n"]
    qq = R.synthetic_qq([8, 8, 8], seed=21)
    h = qe.HPsi()
    h.setup_from(d, inp)
    h.set_qq(qq)
    psi = inp["psi"].copy()
    psi[:, n:] = 0
    psi_t = torch.from_numpy(psi).cuda()
    hpsi_t, spsi_t = torch.empty_like(psi_t), torch.empty_like(psi_t)
    h.h_psi(lda, n, nb, psi_t, hpsi_t)
    h.s_psi(lda, n, nb, psi_t, spsi_t)
    torch.cuda.synchronize()
    href, becp = R.h_psi(d, lda, n, nb, psi, inp["vrs"], inp["vkb"], inp["deeq"], inp["nh_at"], inp["ofsbeta"])
    sref = R.s_psi(d, lda, n, nb, psi, becp, inp["vkb"], qq, inp["nh_at"], inp["ofsbeta"])

    def proj(a, b):  # reference convention: C(i,j) = <a_i|b_j>, stored Fortran (ka, mb) = numpy (mb, ka)
        if d.gamma_only:
            c = 2.0 * (b[:, :n].real @ a[:, :n].real.T + b[:, :n].imag @ a[:, :n].imag.T)
            if d.gstart == 2:
                c -= np.outer(b[:, 0].real, a[:, 0].real)
            return c
        return b[:, :n] @ np.conj(a[:, :n]).T

    dt = np.float64 if d.gamma_only else np.complex128
    nvecx = nb + 2  # leading dimension of the reduced matrices, as in cegterg (hc(nvecx,nvecx))
    hc = np.zeros((nb, nvecx), dtype=dt)
    sc = np.zeros((nb, nvecx), dtype=dt)
    h.dav_gemm_cn(n, psi_t, lda, nb, hpsi_t, lda, nb, hc, nvecx)
    h.dav_gemm_cn(n, psi_t, lda, nb, spsi_t, lda, nb, sc, nvecx)
    assert relerr(hc[:, :nb], proj(psi, href)) < TOL
    assert relerr(sc[:, :nb], proj(psi, sref)) < TOL
    assert np.all(hc[:, nb:] == 0)
    # Rayleigh-Ritz on the host (cdiaghg / rdiaghg are LAPACK on small matrices), basis update on the device
    import scipy.linalg as sl

    Hm, Sm = hc[:, :nb].T, sc[:, :nb].T  # numpy (mb,ka) -> matrix [i,j]
    Hm, Sm = 0.5 * (Hm + Hm.conj().T), 0.5 * (Sm + Sm.conj().T)
    ew, vc = sl.eigh(Hm, Sm)
    vcF = np.ascontiguousarray(vc[:, :notcnv].T)  # Fortran vc(nb, notcnv) memory
    new_t = torch.full((notcnv, lda, 2), float("nan"), device="cuda", dtype=torch.float64)
    new_t[:, n:] = 7.0
    h.dav_gemm_nn(n, spsi_t, lda, nb, vcF, nb, notcnv, new_t, lda)
    torch.cuda.synchronize()
    got = torch.view_as_complex(new_t).cpu().numpy()
    want = vc[:, :notcnv].T @ sref[:, :n]
    assert relerr(got[:, :n], want) < TOL
    assert np.all(got[:, n:] == 7.0 + 7.0j)  # rows beyond n untouched
    # psi_corr = hpsi vc - ew spsi vc (cegterg.f90:353-395): accumulate form
    corr_t = new_t.clone()
    corr_t[:, :n] *= torch.from_numpy(-ew[:notcnv]).cuda()[:, None, None]
    h.dav_gemm_nn(n, hpsi_t, lda, nb, vcF, nb, notcnv, corr_t, lda, accumulate=True)
    torch.cuda.synchronize()
    wantc = vc[:, :notcnv].T @ href[:, :n] - ew[:notcnv, None] * want
    gotc = torch.view_as_complex(corr_t).cpu().numpy()
    assert np.abs(gotc[:, :n] - wantc).max() < 1e-11 * max(np.abs(href).max(), 1.0)
    # g_psi + norms
    rng = np.random.default_rng(3)
    h_diag, s_diag = rng.uniform(0.5, 30.0, n), rng.uniform(0.8, 1.2, n)
    h.set_g_psi_diag(h_diag, s_diag)
    before = gotc.copy()
    h.g_psi(lda, n, notcnv, 1, corr_t, ew[:notcnv].copy())
    torch.cuda.synchronize()
    x = h_diag[None, :] - ew[:notcnv, None] * s_diag[None, :]
    denm = 0.5 * (1.0 + x + np.sqrt(1.0 + (x - 1.0) * (x - 1.0)))
    after = torch.view_as_complex(corr_t).cpu().numpy()
    assert relerr(after[:, :n], before[:, :n] / denm) < 1e-14
    nrm = h.col_dot(n, notcnv, corr_t, lda, corr_t, lda)
    a = after[:, :n]
    ref_n = (np.abs(a) ** 2).sum(1)
    if d.gamma_only:
        ref_n = 2.0 * ref_n - (a[:, 0].real ** 2 if d.gstart == 2 else 0.0)
    assert relerr(nrm, ref_n) < 1e-13
    h.close()


@pytest.mark.parametrize("name", SMALL + ["si2", "al4"])
@pytest.mark.parametrize("m,batch", [(1, 1), (5, 2), (8, 3)])
def test_sum_band_rho(mods, name, m, batch):
    """density part of sum_band (PW/src/sum_band.f90:326-647): rho += sum_b w_b |psi_b(r)|^2; host and device arrays;
    the integral of rho equals sum_b w_b <psi_b|psi_b> nnr (Parseval, unscaled G->R transform)."""
    qe, D, R, torch = mods
    d = D.config_descriptor(name)
    inp = R.synthetic_inputs(d, m, [4, 4], seed=m, poison=True)
    lda, n = inp["lda"], inp["n"]
    rng = np.random.default_rng(m)
    w = rng.uniform(0.0, 2.0, m)
    rho0 = rng.uniform(0.0, 1.0, d.nnr)
    ref = R.sum_band_rho(d, lda, n, m, inp["psi"], w, rho0.copy())
    h = qe.HPsi(batch=batch)
    h.setup_from(d, inp)
    rho = rho0.copy()
    h.sum_band_rho(lda, n, m, inp["psi"], w, rho)
    assert relerr(rho, ref) < 1e-12
    psi = inp["psi"][:, :n]
    if d.gamma_only:
        nrm = 2.0 * (np.abs(psi) ** 2).sum(1) - (np.abs(psi[:, 0]) ** 2 if d.gstart == 2 else 0.0)
    else:
        nrm = (np.abs(psi) ** 2).sum(1)
    assert abs((rho - rho0).sum() / d.nnr - (w * nrm).sum()) < 1e-11
    psi_t = torch.from_numpy(np.nan_to_num(inp["psi"])).cuda()
    rho_t = torch.from_numpy(rho0.copy()).cuda()
    h.sum_band_rho(lda, n, m, psi_t, torch.from_numpy(w).cuda(), rho_t)
    torch.cuda.synchronize()
    assert relerr(rho_t.cpu().numpy(), ref) < 1e-12
    h.close()


def test_sum_band_rho_ausurf_grid(mods):
    qe, D, R, torch = mods
    d = D.config_descriptor("ausurf112")
    m = 3
    inp = R.synthetic_inputs(d, m, [4], seed=1, lda=d.npw, poison=False)
    w = np.array([2.0, 1.5, 0.25])
    ref = R.sum_band_rho(d, d.npw, d.npw, m, inp["psi"], w, np.zeros(d.nnr))
    h = qe.HPsi()
    h.setup_from(d, inp)
    rho = np.zeros(d.nnr)
    h.sum_band_rho(d.npw, d.npw, m, inp["psi"], w, rho)
    assert relerr(rho, ref) < 1e-12
    h.close()


@pytest.mark.parametrize("name", ["si2_small", "si2_small_gamma"])
def test_edge_cases_empty_and_tight(mods, name):
    """m = 0 is a no-op for every entry point (the reference's loops simply do not execute); lda == n (no pad rows);
    a batch larger than the number of bands; a single band on the Gamma path (odd remainder, vloc_psi_acc.f90:93-99)."""
    qe, D, R, torch = mods
    d = D.config_descriptor(name)
    m = 3
    inp = R.synthetic_inputs(d, m, [4, 4], seed=8, lda=d.npw, poison=False)
    lda, n = inp["lda"], inp["n"]
    assert lda == n
    h = qe.HPsi(batch=16)
    h.setup_from(d, inp)
    sentinel = np.full((m, lda), 3.0 + 4.0j)
    out = sentinel.copy()
    h.h_psi(lda, n, 0, inp["psi"], out)
    h.s_psi(lda, n, 0, inp["psi"], out)
    (h.vloc_psi_gamma if d.gamma_only else h.vloc_psi_k)(lda, n, 0, inp["psi"], None, out)
    rho = np.ones(d.nnr)
    h.sum_band_rho(lda, n, 0, inp["psi"], np.zeros(1), rho)
    assert np.array_equal(out, sentinel) and np.all(rho == 1.0)
    ref, _ = R.h_psi(d, lda, n, m, inp["psi"], inp["vrs"], inp["vkb"], inp["deeq"], inp["nh_at"], inp["ofsbeta"])
    h.h_psi(lda, n, m, inp["psi"], out)
    assert relerr(out, ref) < TOL
    one = np.zeros((1, lda), dtype=complex)
    h.h_psi(lda, n, 1, inp["psi"][:1].copy(), one)
    assert relerr(one, ref[:1]) < TOL
    h.close()


@pytest.mark.parametrize("name", ["ortho_small", "si2"])
def test_projectors_3m_adversarial_scales(mods, name):
    """The complex projector GEMMs run in the 3-multiplication form (gemm.cu k_calbec3m / k_addvus3m), whose error bound
    is normwise. Adversarial operands: band columns scaled over 12 decades (1e-6 .. 1e+6) and projectors with
    |Re vkb| = 1e6 |Im vkb|; every band column and every becp column must still meet 1e-11 relative to ITS OWN scale
    (per-column relerr), for calbec, add_vuspsi and the whole h_psi."""
    qe, D, R, torch = mods
    d = D.config_descriptor(name)
    c = d.extra["config"]
    nh = c.get("nh_list") or [c["nh"]] * c["nat"]
    m = 13
    inp = R.synthetic_inputs(d, m, nh, seed=31, poison=False, symmetric=False)
    lda, n = inp["lda"], inp["n"]
    scale = 10.0 ** np.linspace(-6.0, 6.0, m)
    psi = inp["psi"] * scale[:, None]
    vkb = inp["vkb"].copy()
    vkb = vkb.real * 1.0e3 + 1j * vkb.imag * 1.0e-3  # |Re| = 1e6 |Im|
    inp["vkb"] = vkb
    h = qe.HPsi()
    h.setup_from(d, inp)
    becp_ref = R.calbec_k(n, vkb, psi, m)
    becp = np.zeros_like(becp_ref)
    h.calbec(n, None, psi, becp, m, lda=lda)
    assert relerr(becp, becp_ref) < TOL
    ps = R.apply_deeq(becp_ref, inp["deeq"], inp["nh_at"], inp["ofsbeta"])
    href = np.zeros((m, lda), complex)
    R.add_vuspsi_k(n, m, vkb, ps, href)
    hout = np.zeros((m, lda), complex)
    h.add_vuspsi(lda, n, m, hout, becp_ref.copy())
    assert relerr(hout, href) < TOL
    ref, _ = R.h_psi(d, lda, n, m, psi, inp["vrs"], vkb, inp["deeq"], inp["nh_at"], inp["ofsbeta"])
    out = np.full((m, lda), np.nan + 0j)
    h.h_psi(lda, n, m, psi, out)
    assert relerr(out, ref) < TOL
    # the imaginary parts of the projections, which the 3M form obtains by cancellation (P3 + P1 - P2), on their own scale:
    # documented bound eps * |Re| / |Im| ~ 1e-16 * 1e6
    im_err = np.abs(becp.imag - becp_ref.imag).max(axis=1) / np.abs(becp_ref.imag).max(axis=1)
    assert im_err.max() < 1e-8
    h.close()


def test_band_groups_leave_local_becp(mods):
    """The shim's band-group sequence (fortran/qeb200_shim.f90 h_psi_qeb): qeb_h_psi_bgrp on this group's column block,
    then qeb_get_becp for the LOCAL block only - the reference's h_psi_ leaves becp for m_end-m_start+1 columns
    (PW/src/h_psi.f90:57-62); asking for all m columns is an error."""
    qe, D, R, torch = mods
    d = D.config_descriptor("ortho_small")
    m, nb = 7, 3
    inp = R.synthetic_inputs(d, m, [8, 8, 8], seed=4, poison=False)
    lda, n = inp["lda"], inp["n"]
    _, becp_ref = R.h_psi(d, lda, n, m, inp["psi"], inp["vrs"], inp["vkb"], inp["deeq"], inp["nh_at"], inp["ofsbeta"])
    h = qe.HPsi()
    h.setup_from(d, inp)
    out = np.zeros((m, lda), dtype=complex)
    for g in range(nb):
        s, e = qe.divide_all(m, nb, g)
        h.h_psi_bgrp(lda, n, m, inp["psi"], out, nb, g)
        assert relerr(h.get_becp(e - s), becp_ref[s:e]) < TOL
        with pytest.raises(qe.QebError):
            h.get_becp(m)
    h.close()


def test_pin_host_explicit(mods):
    """Page-locking of caller arrays is explicit (qeb_pin_host / qeb_unpin_host): pinned and pageable arrays give the
    same result through the host pipeline; a freed-and-reallocated array is never assumed to be pinned."""
    qe, D, R, torch = mods
    d = D.config_descriptor("ortho_small")
    m = 12
    inp = R.synthetic_inputs(d, m, [8, 8, 8], seed=12)
    lda, n = inp["lda"], inp["n"]
    ref, _ = R.h_psi(d, lda, n, m, inp["psi"], inp["vrs"], inp["vkb"], inp["deeq"], inp["nh_at"], inp["ofsbeta"])
    h = qe.HPsi()
    h.setup_from(d, inp)
    h.set_pipeline(4, 2)
    psi = inp["psi"].copy()
    out = np.full((m, lda), np.nan + 0j)
    h.pin_host(psi)
    h.pin_host(out)
    with pytest.raises(qe.QebError):
        h.pin_host(out)  # already pinned
    h.h_psi(lda, n, m, psi, out)
    assert relerr(out, ref) < TOL
    h.unpin_host(psi)
    h.unpin_host(out)
    with pytest.raises(qe.QebError):
        h.unpin_host(out)
    del psi, out
    for _ in range(3):  # fresh pageable arrays (possibly at the old addresses): no stale registration is used
        p2 = inp["psi"].copy()
        o2 = np.full((m, lda), np.nan + 0j)
        h.h_psi(lda, n, m, p2, o2)
        assert relerr(o2, ref) < TOL
        del p2, o2
    h.close()
