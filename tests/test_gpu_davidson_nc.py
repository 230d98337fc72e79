"""GPU parity of qeb_cegterg with npol = 2 (noncollinear spinors: kdim = kdmx = npwx*npol, cegterg.f90:131-142; g_psi with
h_diag(npwx, npol), g_psi.f90:44-62) against the numpy restatement (oracle/davidson_ref.py, npol = 2) on identical inputs
and against a dense generalized eigensolve of the oracle's spinor H and S."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mods():
    import torch

    if not torch.cuda.is_available():
        pytest.fail("GPU tests need a CUDA device; there is no CPU fallback")
    import qe_b200
    from oracle import davidson_ref as DR

    return qe_b200, DR, torch


def _ctx(qe, c):
    h = qe.HPsi()
    h.setup_from(c["d"], c["inp"])
    h.set_vloc_nc(c["v4"])
    h.set_nonlocal_nc(c["dnc"])
    h.set_qq_nc(c["qnc"])
    h.set_g_psi_diag(c["h_diag"], c["s_diag"])  # (npol, npwx) memory = Fortran h_diag(npwx, npol)
    return h


def test_g_psi_npol2(mods):
    from davidson_case import build_nc

    qe, DR, torch = mods
    c = build_nc("si2_small", 3)
    n, lda = c["n"], c["lda"]
    h = _ctx(qe, c)
    rng = np.random.default_rng(1)
    psi = rng.standard_normal((3, 2 * lda)) + 1j * rng.standard_normal((3, 2 * lda))
    e = np.array([-0.3, 0.1, 0.7])
    ref = psi.copy()
    DR.g_psi_nc(ref, e, n, lda, c["h_diag"], c["s_diag"])
    t = torch.from_numpy(psi.copy()).cuda()
    h.g_psi(lda, n, 3, 2, t, e)
    torch.cuda.synchronize()
    got = t.cpu().numpy()
    assert np.abs(got - ref).max() < 1e-13 * np.abs(ref).max()
    assert np.array_equal(got[:, n:lda], psi[:, n:lda])  # rows beyond n are not touched
    h.close()


@pytest.mark.parametrize("name,nvec,david,nh", [("si2_small", 4, 4, (4, 4)), ("ortho_small", 5, 2, (4, 4, 4)), ("si2", 6, 3, (4, 4))])
@pytest.mark.parametrize("device_evc", [False, True])
def test_cegterg_npol2_matches_restatement(mods, name, nvec, david, nh, device_evc):
    from davidson_case import build_nc, dense_eigenvalues_nc

    qe, DR, torch = mods
    c = build_nc(name, nvec, nh=nh)
    n, lda = c["n"], c["lda"]
    nvecx = david * nvec
    btype = np.ones(nvec, dtype=np.int32)
    btype[-1] = 0
    ethr = 1e-9
    ref_evc = c["evc"].copy()
    e_ref, ref_evc, notcnv_ref, it_ref, nh_ref = DR.davidson(c["h_op"], c["s_op"], c["g_op"], n, lda, nvec, nvecx, ref_evc, ethr,
                                                             True, btype, npol=2)
    h = _ctx(qe, c)
    evc = c["evc"].copy()
    arg = torch.from_numpy(evc).cuda() if device_evc else evc
    e, notcnv, it, nhpsi = h.cegterg(n, lda, nvec, nvecx, arg, ethr, True, btype, npol=2)
    got = arg.cpu().numpy() if device_evc else evc
    scale = max(1.0, np.abs(e_ref).max())
    assert np.abs(e - e_ref).max() < 1e-9 * scale, (e, e_ref)
    assert (notcnv, it, nhpsi) == (notcnv_ref, it_ref, nh_ref)
    assert np.all(got[:, n:lda] == 0) and np.all(got[:, lda + n:] == 0)
    r = c["h_op"](got) - e[:, None] * c["s_op"](got)
    rr = c["h_op"](ref_evc) - e_ref[:, None] * c["s_op"](ref_evc)
    assert np.abs(r).max() < 10 * max(np.abs(rr).max(), 1e-9)
    for ntry in range(3):  # the caller's retry loop with lrot
        if notcnv == 0:
            break
        e, notcnv, it, nhpsi = h.cegterg(n, lda, nvec, nvecx, arg, ethr, True, btype, lrot=True, npol=2)
    assert notcnv == 0
    if not device_evc:  # the dense solve is the slow part of this test; once per case
        w = dense_eigenvalues_nc(c, nvec)
        assert np.abs(e[:-1] - w[:-1]).max() < 1e-7 * scale
        assert abs(e[-1] - w[-1]) < 1e-3
    h.close()


def test_cegterg_npol2_needs_spinor_diagonals(mods):
    from davidson_case import build_nc

    qe, DR, torch = mods
    c = build_nc("si2_small", 4)
    n, lda = c["n"], c["lda"]
    h = _ctx(qe, c)
    h.set_g_psi_diag(c["h_diag"][0], c["s_diag"][0])  # scalar diagonals only
    with pytest.raises(qe.QebError):
        h.cegterg(n, lda, 4, 16, c["evc"].copy(), 1e-8, True, np.ones(4, dtype=np.int32), npol=2)
    h.close()
