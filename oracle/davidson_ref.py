"""TEST INFRASTRUCTURE (oracle): numpy restatement of the block-Davidson solver

    cegterg   KS_Solvers/Davidson/cegterg.f90:20-700   (k-points: complex reduced matrices hc, sc, vc)
    regterg   KS_Solvers/Davidson/regterg.f90:20-640   (Gamma: real hr, sr, vr; 2*dot - G=0 term; Im psi(G=0) = 0)

for one band group and serial diagonalisation (`diaghg` -> LAPACK zhegvx/dsygvx; here scipy.linalg.eigh on the same
matrices). The operators are passed in as callables with the reference's EXTERNAL signatures (h_psi_ptr, s_psi_ptr,
g_psi_ptr, cegterg.f90:110-118). Arrays are numpy (columns, npwx) = Fortran (npwx, columns). Only tests/ imports this.

Parity status: unpinned by reference fixtures (pw.x cannot be built here); the restatement follows the cited lines and is
checked against a dense generalized eigensolve of the same operators (tests/test_oracle.py)."""
from __future__ import annotations

import numpy as np
import scipy.linalg as sl

MAXTER = 20  # cegterg.f90:71


def _proj(a, b, n, gamma, gstart):
    """C[i, j] = <a_i | b_j> over n plane waves (ZGEMM 'C','N', cegterg.f90:209; Gamma: 2*DGEMM('T','N') - DGER at G=0,
    regterg.f90:204-205)"""
    if gamma:
        c = 2.0 * (a[:, :n].real @ b[:, :n].real.T + a[:, :n].imag @ b[:, :n].imag.T)
        if gstart == 2:
            c -= np.outer(a[:, 0].real, b[:, 0].real)
        return c
    return np.conj(a[:, :n]) @ b[:, :n].T


def diaghg(n, m, h, s):
    """lowest m eigenpairs of h v = e s v on the leading n x n blocks, v^H s v = 1 (LAXlib cdiaghg/rdiaghg)"""
    ew, v = sl.eigh(h[:n, :n], s[:n, :n], subset_by_index=[0, m - 1])
    return ew, v


def g_psi(psi, e, n, h_diag, s_diag):
    """PW/src/g_psi.f90:10-70, the TEST_NEW_PRECONDITIONING branch the file defines"""
    x = h_diag[None, :n] - e[:, None] * s_diag[None, :n]
    denm = 0.5 * (1.0 + x + np.sqrt(1.0 + (x - 1.0) * (x - 1.0)))
    psi[:, :n] /= denm


def g_psi_nc(psi, e, n, lda, h_diag, s_diag):
    """g_psi with npol = 2 (PW/src/g_psi.f90:44-62): psi (m, 2*lda), h_diag / s_diag (2, lda) = Fortran (npwx, npol)"""
    for ipol in range(2):
        x = h_diag[ipol][None, :n] - e[:, None] * s_diag[ipol][None, :n]
        denm = 0.5 * (1.0 + x + np.sqrt(1.0 + (x - 1.0) * (x - 1.0)))
        psi[:, ipol * lda:ipol * lda + n] /= denm


def davidson(h_op, s_op, g_op, npw, npwx, nvec, nvecx, evc, ethr, uspp, btype, lrot=False, gamma=False, gstart=1,
             maxter=MAXTER, npol=1):
    """returns (e, evc, notcnv, dav_iter, nhpsi); evc is also updated in place.

    h_op(psi) -> hpsi, s_op(psi) -> spsi on (m, npwx*npol) blocks; g_op(psi, e) preconditions in place.
    npol = 2 (noncollinear spinors): kdim = kdmx = npwx*npol, i.e. the products run over the pad rows too
    (cegterg.f90:131-142)."""
    if nvec > nvecx // 2:
        raise ValueError("nvecx is too small")  # cegterg.f90:125
    empty_ethr = max(ethr * 5.0, 1e-5)  # :128
    dt = np.float64 if gamma else np.complex128
    npw1 = npw  # plane waves of one component
    kdmx = npwx * npol
    if npol != 1:
        npw = kdmx  # kdim
    psi = np.zeros((nvecx, kdmx), complex)
    hpsi = np.zeros((nvecx, kdmx), complex)
    spsi = np.zeros((nvecx, kdmx), complex) if uspp else None
    hc = np.zeros((nvecx, nvecx), dt)
    sc = np.zeros((nvecx, nvecx), dt)
    vc = np.zeros((nvecx, nvecx), dt)
    ew = np.zeros(nvecx)
    e = np.zeros(nvec)
    conv = np.zeros(nvec, bool)
    notcnv, nbase, nhpsi, dav_iter = nvec, nvec, 0, 0
    psi[:nvec] = evc[:nvec]
    if gamma and gstart == 2:
        psi[:nvec, 0] = psi[:nvec, 0].real  # regterg.f90:174
    hpsi[:nvec] = h_op(psi[:nvec])
    nhpsi += nvec
    if uspp:
        spsi[:nvec] = s_op(psi[:nvec])
    sb = spsi if uspp else psi
    hc[:nbase, :nbase] = _proj(psi[:nbase], hpsi[:nbase], npw, gamma, gstart)  # :209-229
    sc[:nbase, :nbase] = _proj(psi[:nbase], sb[:nbase], npw, gamma, gstart)

    def symmetrise(lo, hi):  # :241-254, :520-533: lower triangle is the master, diagonal strictly real
        for n in range(hi):
            if n >= lo:
                hc[n, n] = hc[n, n].real
                sc[n, n] = sc[n, n].real
            for m in range(max(n + 1, lo), hi):
                hc[n, m] = np.conj(hc[m, n])
                sc[n, m] = np.conj(sc[m, n])

    symmetrise(0, nbase)
    if lrot:  # :257-267
        vc[:nbase, :nbase] = np.eye(nbase)
        e[:] = np.diag(hc)[:nvec].real
    else:
        w, v = diaghg(nbase, nvec, hc, sc)  # :271-285
        ew[:nvec] = w
        vc[:nbase, :nvec] = v
        e[:] = w
    for kter in range(1, maxter + 1):
        dav_iter = kter
        np_ = 0
        for n in range(nvec):  # :305-334: coefficients of the unconverged roots first
            if not conv[n]:
                if np_ != n:
                    vc[:, np_] = vc[:, n]
                ew[nbase + np_] = e[n]
                np_ += 1
        nb1 = nbase
        new = slice(nb1, nb1 + notcnv)
        # ( H - e S ) |psi> of the unconverged roots (:340-395)
        psi[new, :npw] = vc[:nbase, :notcnv].T @ sb[:nbase, :npw]
        psi[new, :npw] *= -ew[new, None]
        psi[new, :npw] += vc[:nbase, :notcnv].T @ hpsi[:nbase, :npw]
        psi[new, npw1:npwx] = 0.0  # :400
        if npol == 2:
            psi[new, npwx + npw1:] = 0.0  # :401
        g_op(psi[new], ew[new])  # :407
        # normalise the correction vectors (:415-460; regterg.f90:358-361); npol = 2: both components (:428-434), whose
        # pad rows are zero here
        nrm = np.einsum("ij,ij->i", np.conj(psi[new, :npw]), psi[new, :npw]).real
        if gamma:
            nrm = 2.0 * nrm
            if gstart == 2:
                nrm -= psi[new, 0].real * psi[new, 0].real
        psi[new, :npw] /= np.sqrt(nrm)[:, None]
        if gamma and gstart == 2:
            psi[new, 0] = psi[new, 0].real  # regterg.f90:375
        hpsi[new] = h_op(psi[new])  # :465
        nhpsi += notcnv
        if uspp:
            spsi[new] = s_op(psi[new])
        # update the reduced matrices: rows nb1.. (:478-499), then the Hermitian fill
        tot = nbase + notcnv
        hc[nb1:tot, :tot] = _proj(hpsi[new], psi[:tot], npw, gamma, gstart)
        sc[nb1:tot, :tot] = _proj(sb[new], psi[:tot], npw, gamma, gstart)
        nbase = tot
        symmetrise(nb1, nbase)
        w, v = diaghg(nbase, nvec, hc, sc)  # :545
        ew[:nvec] = w
        vc[:nbase, :nvec] = v
        thr = np.where(np.asarray(btype[:nvec]) == 1, ethr, empty_ethr)  # :556-562
        conv = np.abs(ew[:nvec] - e) < thr
        notcnv = int(np.count_nonzero(~conv))
        e[:] = ew[:nvec]
        if notcnv == 0 or nbase + notcnv > nvecx or dav_iter == maxter:  # :576-669
            evc[:nvec, :npw] = vc[:nbase, :nvec].T @ psi[:nbase, :npw]  # rows npw+1..npwx of evc are not written (:588)
            if notcnv == 0 or dav_iter == maxter:
                break
            psi[:nvec] = evc[:nvec]
            if uspp:
                t = vc[:nbase, :nvec].T @ spsi[:nbase, :npw]
                spsi[:nvec, :npw] = t
            t = vc[:nbase, :nvec].T @ hpsi[:nbase, :npw]
            hpsi[:nvec, :npw] = t
            nbase = nvec
            hc[:nbase, :nbase] = np.diag(e)
            sc[:nbase, :nbase] = np.eye(nbase)
            vc[:nbase, :nbase] = np.eye(nbase)
    return e, evc, notcnv, dav_iter, nhpsi
