"""CPU tests of the oracle itself: the restatement must reproduce every
known-answer fact the reference tree holds for this path."""
import json
import os

import numpy as np
import pytest

from oracle import descriptor as D
from oracle import hpsi_ref as R


@pytest.fixture(scope="module")
def kat(golden_dir):
    return json.load(open(os.path.join(golden_dir, "fft_test0.json")))


@pytest.fixture(scope="module")
def kat_desc(kat):
    A = np.array(kat["at_rows"]) * kat["alat_in"]
    alat = np.sqrt((A[:, 0] ** 2).sum())
    return D.build_descriptor(A / alat, alat, kat["ecutwfc"], gamma_only=False), alat


def test_descriptor_matches_fft_test0_header(kat, kat_desc):
    d, alat = kat_desc
    assert abs(alat - kat["alat"]) < 1e-6  # the reference literal 18.65 is single precision
    assert [d.nr1, d.nr2, d.nr3] == kat["grid"]
    assert d.ngm == kat["g_smooth"]
    assert d.npw == kat["g_pw"]
    assert d.nst == kat["sticks_pw"]


def test_descriptor_matches_testsuite_si():
    # test-suite/pw_scf/benchmark.out.git.inp=scf.in: FFT 15x15x15, 180 / 186 PWs
    d = D.config_descriptor("si2_small")
    assert (d.nr1, d.nr2, d.nr3) == (15, 15, 15)
    assert d.npw in (180, 186)
    # SURVEY 8 table
    d = D.config_descriptor("si2")
    assert (d.nr1, d.nr2, d.nr3, d.npw) == (25, 25, 25, 734)
    d = D.config_descriptor("al4")
    assert (d.nr1, d.nr2, d.nr3, d.npw) == (27, 27, 27, 1246)


def test_good_fft_order():
    assert [D.good_fft_order(n) for n in (119, 167, 7, 11, 13, 121, 128)] == [120, 180, 8, 12, 15, 125, 128]


def test_hpsort_eps_ties_by_index():
    ra = np.array([3.0, 1.0, 1.0 + 1e-10, 2.0, 1.0 - 1e-10])
    s, ind = D.hpsort_eps(ra)
    assert list(ind) == [2, 3, 5, 4, 1]
    rng = np.random.default_rng(0)
    x = rng.random(1000)
    s, ind = D.hpsort_eps(x)
    assert np.all(np.diff(s) >= 0) and np.array_equal(x[ind - 1], s)


def test_g_sorted_and_maps(kat_desc):
    d, _ = kat_desc
    assert np.all(np.diff(d.gg) > -1e-8)
    assert d.gstart == 2 and d.nl[0] == 1
    # nl is a bijection onto distinct box points, igk sorted by |k+G|^2
    assert np.unique(d.nl).size == d.ngm
    assert np.all(np.diff(d.g2kin) > -1e-6)
    # every PW lies on a wave stick and active plane
    col = (d.nl[d.igk - 1] - 1) % (d.nr1 * d.nr2)
    assert np.all(d.isind[col] == 1)
    assert np.all(d.iplw[col % d.nr1] == 1)


def _kat_box(kat):
    nr = kat["grid"]
    box = np.zeros(nr[0] * nr[1] * nr[2], dtype=np.complex128)
    for key, (re_, im_) in kat["coeff"].items():
        i, j, k = (int(t) - 1 for t in key.split(","))
        box[i + j * nr[0] + k * nr[0] * nr[1]] = re_ + 1j * im_
    return nr, box


def test_invfft_known_answer_numpy(kat, kat_desc):
    d, _ = kat_desc
    nr, box = _kat_box(kat)
    out = R.cfft3ds(box, nr, +1, d.isind, d.iplw).reshape(nr[2], nr[1], nr[0])
    gold = np.array(kat["real_space_5x5x5_kji"]).reshape(5, 5, 5, 2)
    got = out[:5, :5, :5]
    assert np.abs(got.real - gold[..., 0]).max() < 1.01e-6
    assert np.abs(got.imag - gold[..., 1]).max() < 1.01e-6
    # analytic form, full precision
    z, y, x = np.meshgrid(np.arange(5), np.arange(5), np.arange(5), indexing="ij")
    ana = 1 + 0.5j * np.exp(2j * np.pi * x / nr[0]) + 0.3j * np.exp(2j * np.pi * y / nr[1]) + 0.7j * np.exp(2j * np.pi * z / nr[2])
    assert np.abs(got - ana).max() < 1e-13
    # R->G returns the inputs (scaling 1/N)
    back = R.cfft3ds(out.reshape(-1), nr, -1, d.isind, d.iplw)
    sel = np.nonzero(box)[0]
    assert np.abs(back[sel] - box[sel]).max() < 1e-14


@pytest.mark.skipif(R.ref_lib() is None, reason="oracle/_ref not built")
def test_invfft_known_answer_reference_engine(kat, kat_desc):
    d, _ = kat_desc
    nr, box = _kat_box(kat)
    out = R.ref_cfft3ds(box, nr, +1, d.isind, d.iplw).reshape(nr[2], nr[1], nr[0])
    gold = np.array(kat["real_space_5x5x5_kji"]).reshape(5, 5, 5, 2)
    assert np.abs(out[:5, :5, :5].real - gold[..., 0]).max() < 1.01e-6
    assert np.abs(out[:5, :5, :5].imag - gold[..., 1]).max() < 1.01e-6


@pytest.mark.skipif(R.ref_lib() is None, reason="oracle/_ref not built")
@pytest.mark.parametrize("name", ["si2_small", "ortho_small", "si2_small_gamma", "ortho_small_gamma"])
def test_restatement_vs_reference_fft_engine(name):
    """numpy restatement of vloc_psi == restated driver on the reference's C FFT."""
    d = D.config_descriptor(name)
    c = d.extra["config"]
    m = c["nbnd"]
    inp = R.synthetic_inputs(d, m, [c["nh"]] * c["nat"], seed=3, poison=False)
    lda, n = inp["lda"], inp["n"]
    h0 = np.zeros((m, lda), dtype=np.complex128)
    h1 = np.zeros((m, lda), dtype=np.complex128)
    (R.vloc_psi_gamma if d.gamma_only else R.vloc_psi_k)(d, lda, n, m, inp["psi"], inp["vrs"], h0)
    R.ref_vloc_psi(d, lda, n, m, inp["psi"], inp["vrs"], h1)
    scale = np.abs(h0[:, :n]).max()
    assert np.abs(h0[:, :n] - h1[:, :n]).max() <= 1e-13 * max(scale, 1.0)
    assert np.all(h1[:, n:] == 0)


def test_random_box_fft_against_full_fft():
    """cfft3ds on a contract-respecting box (zeros off the sticks) == plain 3-D FFT."""
    d = D.config_descriptor("ortho_small")
    nr = (d.nr1, d.nr2, d.nr3)
    rng = np.random.default_rng(5)
    box = np.zeros(d.nnr, dtype=np.complex128)
    idx = d.nl[d.igk - 1] - 1
    box[idx] = rng.standard_normal(idx.size) + 1j * rng.standard_normal(idx.size)
    a = R.cfft3ds(box, nr, +1, d.isind, d.iplw).reshape(nr[2], nr[1], nr[0])
    b = np.fft.ifftn(box.reshape(nr[2], nr[1], nr[0])) * d.nnr
    assert np.abs(a - b).max() < 1e-12


def test_hpsi_hermitian_and_gamma_real():
    """H built from the restatement is Hermitian: <a|H b> = conj(<b|H a>)."""
    for name in ("ortho_small", "ortho_small_gamma"):
        d = D.config_descriptor(name)
        c = d.extra["config"]
        inp = R.synthetic_inputs(d, 2, [c["nh"]] * c["nat"], seed=11, poison=False)
        lda, n = inp["lda"], inp["n"]
        h, becp = R.h_psi(d, lda, n, 2, inp["psi"], inp["vrs"], inp["vkb"], inp["deeq"], inp["nh_at"], inp["ofsbeta"])
        a, b = inp["psi"][0, :n], inp["psi"][1, :n]
        if d.gamma_only:
            def dot(x, y):  # real-space inner product of Gamma-trick vectors
                return 2.0 * np.vdot(x, y).real - (x[0].conjugate() * y[0]).real
            assert abs(dot(a, h[1, :n]) - dot(b, h[0, :n])) < 1e-12
            assert np.all(h[:, 0].imag == 0)
        else:
            assert abs(np.vdot(a, h[1, :n]) - np.conj(np.vdot(b, h[0, :n]))) < 1e-12
        assert np.all(h[:, n:] == 0)


def test_divide_all():
    c, dsp = R.divide_all(10, 4)
    assert list(c) == [3, 3, 2, 2] and list(dsp) == [0, 3, 6, 8]


def test_s_psi_restatement_hermitian_and_copy():
    """S = 1 + sum |beta_i> qq_ij <beta_j| (PW/src/s_psi.f90:10-262) is Hermitian for symmetric qq; without qq
    (norm-conserving, :120) it is the copy of all lda rows."""
    for name in ("ortho_small", "ortho_small_gamma"):
        d = D.config_descriptor(name)
        c = d.extra["config"]
        nh = [c["nh"]] * c["nat"]
        inp = R.synthetic_inputs(d, 2, nh, seed=4, poison=False)
        lda, n = inp["lda"], inp["n"]
        _, becp = R.h_psi(d, lda, n, 2, inp["psi"], inp["vrs"], inp["vkb"], inp["deeq"], inp["nh_at"], inp["ofsbeta"])
        assert np.array_equal(R.s_psi(d, lda, n, 2, inp["psi"], becp), inp["psi"])
        qq = R.synthetic_qq(nh, seed=4)
        s = R.s_psi(d, lda, n, 2, inp["psi"], becp, inp["vkb"], qq, inp["nh_at"], inp["ofsbeta"])
        assert np.array_equal(s[:, n:], inp["psi"][:, n:])
        a, b = inp["psi"][0, :n], inp["psi"][1, :n]
        if d.gamma_only:
            def dot(x, y):
                return 2.0 * np.vdot(x, y).real - (x[0].conjugate() * y[0]).real
            assert abs(dot(a, s[1, :n]) - dot(b, s[0, :n])) < 1e-12
        else:
            assert abs(np.vdot(a, s[1, :n]) - np.conj(np.vdot(b, s[0, :n]))) < 1e-12


def test_ylmr2_restatement_matches_scipy():
    """oracle.uspp_ref.ylmr2 (upflib/ylmr2.f90) against scipy's complex harmonics: m=0 -> Y_l0, m>0 -> sqrt2 Re Y_lm
    (cos slot) and sqrt2 Im Y_lm (sin slot)."""
    from scipy import special

    from oracle import uspp_ref as U

    rng = np.random.default_rng(0)
    g = rng.standard_normal((200, 3))
    g[0] = [0.0, 0.0, 1.3]
    g[1] = [0.0, -0.7, 0.2]
    gg = (g * g).sum(1)
    lmax = 4
    y = U.ylmr2((lmax + 1) ** 2, g, gg)
    theta = np.arccos(g[:, 2] / np.sqrt(gg))
    phi = np.arctan2(g[:, 1], g[:, 0])
    sph = getattr(special, "sph_harm_y", None)
    for l in range(lmax + 1):
        for m in range(l + 1):
            Y = sph(l, m, theta, phi) if sph else special.sph_harm(m, l, phi, theta)
            if m == 0:
                assert np.abs(y[:, l * l] - Y.real).max() < 1e-13
            else:
                assert np.abs(y[:, l * l + 2 * m - 1] - np.sqrt(2.0) * Y.real).max() < 1e-13
                assert np.abs(y[:, l * l + 2 * m] - np.sqrt(2.0) * Y.imag).max() < 1e-13


def test_init_us_2_restatement_translation_phase():
    """moving an atom by a lattice vector leaves vkb unchanged; moving it by tau multiplies by exp(-i(k+G).tau)."""
    from oracle import uspp_ref as U

    d = D.config_descriptor("si2_small")
    us = U.synthetic_uspp(1, 2, seed=1)
    tpiba = 2.0 * np.pi / d.alat
    nr = (d.nr1, d.nr2, d.nr3)
    tau0 = np.array([[0.1, 0.2, -0.05]])
    shift = np.array([0.03, -0.02, 0.07])
    vk = []
    for tau in (tau0, tau0 + d.at[:, 0][None, :], tau0 + shift[None, :]):
        e = U.eigts(d.bg, tau, nr)
        vk.append(U.init_us_2(d.npw, d.npw, d.igk, d.xk, tau, [1], tpiba, nr, e, d.mill, d.g, us))
    assert np.abs(vk[0] - vk[1]).max() < 1e-12 * np.abs(vk[0]).max() * 10
    kg = d.g[d.igk - 1] + d.xk[None, :]
    ph = np.exp(-1j * 2.0 * np.pi * (kg @ shift))
    assert np.abs(vk[2] - vk[0] * ph[None, :]).max() < 1e-12
    assert np.abs(U.g2_kin(d.g, d.igk, d.xk, d.tpiba2) - d.g2kin).max() == 0.0


def test_sum_band_rho_restatement_parseval():
    """oracle sum_band_rho (PW/src/sum_band.f90:326-647): integral of rho = sum_b w_b <psi_b|psi_b> nnr (k and Gamma,
    odd band count exercises the single-band Gamma transform)."""
    for name in ("si2_small", "si2_small_gamma"):
        d = D.config_descriptor(name)
        m = 3
        inp = R.synthetic_inputs(d, m, [4, 4], seed=1)
        n = inp["n"]
        w = np.array([2.0, 1.0, 0.5])
        rho = R.sum_band_rho(d, inp["lda"], n, m, inp["psi"], w, np.zeros(d.nnr))
        assert rho.min() >= 0.0
        psi = inp["psi"][:, :n]
        nrm = (np.abs(psi) ** 2).sum(1)
        if d.gamma_only:
            nrm = 2.0 * nrm - (np.abs(psi[:, 0]) ** 2 if d.gstart == 2 else 0.0)
        assert abs(rho.sum() / d.nnr - (w * nrm).sum()) < 1e-12


def test_sum_band_rho_nc_restatement():
    """oracle sum_band_rho_nc (PW/src/sum_band.f90:555-575, get_rho_domag :699-729) against (a) the Pauli-matrix form
    rho_k(r) = w psi^+(r) sigma_k psi(r) from a dense numpy transform of the two components, and (b) the decomposition
    the library uses (density_nc.cu): four scalar densities A, B, C = |u + d|^2, D = |u - i d|^2."""
    d = D.config_descriptor("ortho_small")
    m = 3
    inp = R.synthetic_inputs(d, 2 * m, [4, 4, 4], seed=5, poison=False)
    lda, n = inp["lda"], inp["n"]
    evc = np.ascontiguousarray(inp["psi"]).reshape(m, 2 * lda)
    w = np.array([2.0, 0.75, 0.1])
    rho0 = np.random.default_rng(3).uniform(-1.0, 1.0, (4, d.nnr))
    rho = R.sum_band_rho_nc(d, lda, n, m, evc, w, rho0.copy())
    idx = d.nl[d.igk[:n] - 1] - 1
    sig = np.array([[[1, 0], [0, 1]], [[0, 1], [1, 0]], [[0, -1j], [1j, 0]], [[1, 0], [0, -1]]])
    ref = rho0.copy()
    for b in range(m):
        sp = []
        for ipol in range(2):
            box = np.zeros(d.nnr, complex)
            box[idx] = evc[b, ipol * lda:ipol * lda + n]
            sp.append(np.fft.ifftn(box.reshape(d.nr3, d.nr2, d.nr1)).ravel() * d.nnr)
        sp = np.array(sp)
        for k in range(4):
            ref[k] += w[b] * np.einsum("ar,ab,br->r", np.conj(sp), sig[k], sp).real
    assert np.abs(rho - ref).max() < 1e-12 * np.abs(ref).max()
    # without magnetisation only the density is accumulated
    r1 = R.sum_band_rho_nc(d, lda, n, m, evc, w, rho0[0].copy(), domag=False)
    assert np.abs(r1 - rho[0]).max() < 1e-13 * np.abs(rho[0]).max()
    # the four-scalar-density decomposition
    z = lambda: np.zeros(d.nnr)
    up, dw = evc[:, :lda], evc[:, lda:]
    A = R.sum_band_rho(d, lda, n, m, up, w, z())
    B = R.sum_band_rho(d, lda, n, m, dw, w, z())
    C = R.sum_band_rho(d, lda, n, m, up + dw, w, z())
    Dd = R.sum_band_rho(d, lda, n, m, up - 1j * dw, w, z())
    dec = rho0 + np.array([A + B, C - (A + B), Dd - (A + B), A - B])
    for k in range(4):
        assert np.abs(dec[k] - rho[k]).max() < 1e-12 * np.abs(rho[k]).max()


def test_vloc_psi_nc_reduces_to_scalar_and_is_hermitian():
    """vloc_psi_nc (vloc_psi_acc.f90:269-369): with v2=v3=v4=0 it is vloc_psi_k on both components; with a general
    real v(:,1:4) the 2x2 operator is Hermitian: <phi|V psi> = conj(<psi|V phi>) on the sphere."""
    d = D.config_descriptor("si2_small")
    m = 2
    inp = R.synthetic_inputs(d, 2 * m, [4, 4], seed=2, poison=False)
    lda, n = inp["lda"], inp["n"]
    psi = np.ascontiguousarray(inp["psi"]).reshape(m, 2 * lda)
    v4 = np.zeros((4, d.nnr))
    v4[0] = inp["vrs"]
    a = R.vloc_psi_nc(d, lda, n, m, psi, v4, np.zeros((m, 2 * lda), complex))
    b = R.vloc_psi_k(d, lda, n, 2 * m, psi.reshape(2 * m, lda), inp["vrs"], np.zeros((2 * m, lda), complex))
    assert np.abs(a.reshape(2 * m, lda) - b).max() < 1e-13
    c = R.vloc_psi_nc(d, lda, n, m, psi, v4, np.zeros((m, 2 * lda), complex), domag=False)
    assert np.abs(c - a).max() < 1e-13
    rng = np.random.default_rng(0)
    v4 = rng.uniform(-1, 1, (4, d.nnr))
    hv = R.vloc_psi_nc(d, lda, n, m, psi, v4, np.zeros((m, 2 * lda), complex))
    live = np.zeros(2 * lda, bool)
    live[:n] = live[lda:lda + n] = True
    x01 = np.vdot(psi[0, live], hv[1, live])
    x10 = np.vdot(psi[1, live], hv[0, live])
    assert abs(x01 - np.conj(x10)) < 1e-12 * max(1.0, abs(x01))


def test_h_psi_nc_block_diagonal_limit():
    """h_psi_nc with v2=v3=v4=0 and deeq_nc(:,:,:,1)=deeq_nc(:,:,:,4)=deeq, (:,:,:,2:3)=0 (what newd sets without
    magnetisation) is the scalar h_psi on 2m columns (h_psi.f90:138-148; add_vuspsi.f90:192-223 vs :151-155)."""
    d = D.config_descriptor("ortho_small")
    c = d.extra["config"]
    nh = [c["nh"]] * c["nat"]
    m = 2
    inp = R.synthetic_inputs(d, 2 * m, nh, seed=8, poison=False)
    lda, n = inp["lda"], inp["n"]
    psi = np.ascontiguousarray(inp["psi"]).reshape(m, 2 * lda)
    v4 = np.zeros((4, d.nnr))
    v4[0] = inp["vrs"]
    nat, nhm = len(nh), max(nh)
    dnc = np.zeros((4, nat, nhm, nhm), complex)
    dnc[0] = dnc[3] = inp["deeq"]
    a, ba = R.h_psi_nc(d, lda, n, m, psi, v4, inp["vkb"], dnc, inp["nh_at"], inp["ofsbeta"])
    b, bb = R.h_psi(d, lda, n, 2 * m, psi.reshape(2 * m, lda), inp["vrs"], inp["vkb"], inp["deeq"], inp["nh_at"],
                    inp["ofsbeta"])
    assert np.abs(a.reshape(2 * m, lda) - b).max() / np.abs(b).max() < 1e-13
    assert np.abs(ba.reshape(2 * m, -1) - bb).max() < 1e-13
