"""CPU checks of the Davidson restatement (oracle/davidson_ref.py, cegterg.f90 / regterg.f90): its eigenvalues against a
dense generalized eigensolve of the same operators, and the bookkeeping the reference guarantees."""
import numpy as np
import pytest

from oracle import davidson_ref as DR

from davidson_case import build, dense_eigenvalues


@pytest.mark.parametrize("name", ["si2_small", "si2_small_gamma"])
def test_davidson_restatement_vs_dense(name):
    nvec = 4
    c = build(name, nvec)
    d, n, lda = c["d"], c["n"], c["lda"]
    evc = c["evc"].copy()
    for ntry in range(3):  # the caller's retry loop (PW/src/c_bands.f90: diag_bands calls cegterg until notconv == 0)
        e, evc, notcnv, it, nhpsi = DR.davidson(c["h_op"], c["s_op"], c["g_op"], n, lda, nvec, 4 * nvec, evc, 1e-11, True,
                                                np.ones(nvec, int), gamma=d.gamma_only, gstart=d.gstart)
        if notcnv == 0:
            break
    w = dense_eigenvalues(c, nvec)
    assert notcnv == 0 and it < 20
    assert np.abs(e - w).max() < 1e-9 * max(1.0, np.abs(w).max())
    assert nhpsi >= nvec + 1
    # eigenvector residuals with the same operators
    r = c["h_op"](evc)[:, :n] - e[:, None] * c["s_op"](evc)[:, :n]
    assert np.abs(r).max() < 1e-5


def test_davidson_restatement_lrot_and_empty_bands():
    """lrot = .TRUE. starts from vc = 1, e = diag(hc) (cegterg.f90:257-267); empty bands converge to max(5 ethr, 1e-5)"""
    nvec = 4
    c = build("si2_small", nvec, seed=3)
    n, lda = c["n"], c["lda"]
    evc = c["evc"].copy()
    btype = np.array([1, 1, 0, 0])
    for ntry in range(4):
        e, evc, notcnv, it, nhpsi = DR.davidson(c["h_op"], c["s_op"], c["g_op"], n, lda, nvec, 2 * nvec, evc, 1e-9, True, btype,
                                                lrot=ntry > 0)
        if notcnv == 0:
            break
    w = dense_eigenvalues(c, nvec)
    assert notcnv == 0
    assert np.abs(e[:2] - w[:2]).max() < 1e-7
    assert np.abs(e - w).max() < 1e-3


def test_davidson_restatement_npol2_vs_dense():
    """npol = 2 (cegterg.f90:131-142: kdim = kdmx = npwx*npol; :401 pad rows of the second component; :428-434 norm over
    both components; g_psi.f90:44-62 with h_diag(npwx, npol)) against a dense eigensolve of the spinor H and S"""
    from davidson_case import build_nc, dense_eigenvalues_nc

    nvec = 4
    c = build_nc("si2_small", nvec)
    n, lda = c["n"], c["lda"]
    evc = c["evc"].copy()
    for ntry in range(4):
        e, evc, notcnv, it, nhpsi = DR.davidson(c["h_op"], c["s_op"], c["g_op"], n, lda, nvec, 4 * nvec, evc, 1e-11, True,
                                                np.ones(nvec, int), npol=2, lrot=ntry > 0)
        if notcnv == 0:
            break
    w = dense_eigenvalues_nc(c, nvec)
    assert notcnv == 0
    assert np.abs(e - w).max() < 1e-9 * max(1.0, np.abs(w).max())
    assert np.all(evc[:, n:lda] == 0) and np.all(evc[:, lda + n:] == 0)
    r = c["h_op"](evc) - e[:, None] * c["s_op"](evc)
    assert np.abs(r).max() < 1e-5
