"""ORACLE (test infrastructure, not product code) — parity at the h_psi boundary
is UNPINNED by the reference's own tests: the reference
holds no golden vector for h_psi(psi)->hpsi, and pw.x cannot be built here (no
Fortran compiler). What IS pinned: the FFT sign/scale/layout conventions against
FFTXlib/examples/reference/fft_test0.out, the descriptor builder against that
file's grid/stick/G counts and the test-suite's Si numbers, and this numpy
restatement against the reference's own C FFT engine (oracle/_ref).

CPU restatement (numpy, FP64) of PWscf's H|psi> application. Follows (paths
relative to /root/reference):
  h_psi_            PW/src/h_psi.f90:83-302   (kinetic :138-147, V_NL :229-236, Gamma fix :294-295)
  vloc_psi_k_acc    PW/src/vloc_psi_acc.f90:148-266
  vloc_psi_gamma_acc PW/src/vloc_psi_acc.f90:9-145
  wave_g2r/wave_r2g Modules/fft_wave.f90:31-159
  fftx_c2psi_k/psi2c_k        FFTXlib/src/fft_helper_subroutines.f90:447-511,751-795
  fftx_c2psi_gamma/psi2c_gamma FFTXlib/src/fft_helper_subroutines.f90:317-409,644-715
  cfft3ds           FFTXlib/src/fft_scalar.FFTW3.f90:498-694 (stick/plane skipping, scaling)
  calbec_k/gamma    Modules/becmod.f90:431-486,274-334
  add_vuspsi_k/gamma PW/src/add_vuspsi.f90:125-171,68-122
  divide_all        UtilXlib/divide.f90:55-106

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import
this module.
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

from .descriptor import Descriptor

_HERE = os.path.dirname(os.path.abspath(__file__))


# ----------------------------------------------------------------------------
# 3-D "reduced" FFT, reference pass order
# ----------------------------------------------------------------------------
def cfft3ds(f: np.ndarray, nr, isign: int, do_fft_z: np.ndarray, do_fft_y: np.ndarray) -> np.ndarray:
    """f: flat complex box (x fastest) of length nr1*nr2*nr3; returns new flat array.
    isign>0: G->R, exp(+i), unscaled; z on masked columns, y on masked x-planes, x on all.
    isign<0: R->G, exp(-i): x all, y masked planes, z masked columns, then *1/N."""
    nr1, nr2, nr3 = nr
    box = np.array(f, dtype=np.complex128).reshape(nr3, nr2, nr1)  # [z][y][x]
    zmask = do_fft_z.reshape(nr2, nr1) > 0  # [y][x]
    ymask = do_fft_y == 1  # [x]
    if isign > 0:
        cols = box[:, zmask]  # (nr3, ncols)
        box[:, zmask] = np.fft.ifft(cols, axis=0) * nr3
        pl = box[:, :, ymask]
        box[:, :, ymask] = np.fft.ifft(pl, axis=1) * nr2
        box = np.fft.ifft(box, axis=2) * nr1
    else:
        box = np.fft.fft(box, axis=2)
        pl = box[:, :, ymask]
        box[:, :, ymask] = np.fft.fft(pl, axis=1)
        cols = box[:, zmask]
        box[:, zmask] = np.fft.fft(cols, axis=0)
        box = box * (1.0 / (nr1 * nr2 * nr3))
    return box.reshape(-1)


# ----------------------------------------------------------------------------
# V_loc
# ----------------------------------------------------------------------------
def vloc_psi_k(d: Descriptor, lda: int, n: int, m: int, psi: np.ndarray, v: np.ndarray, hpsi: np.ndarray):
    """psi, hpsi: (m, lda) complex (row b = Fortran column b). hpsi is INOUT."""
    nr = (d.nr1, d.nr2, d.nr3)
    idx = d.nl[d.igk[:n] - 1] - 1
    for b in range(m):
        box = np.zeros(d.nnr, dtype=np.complex128)
        box[idx] = psi[b, :n]
        box = cfft3ds(box, nr, +1, d.isind, d.iplw)
        box = box * v
        box = cfft3ds(box, nr, -1, d.isind, d.iplw)
        hpsi[b, :n] += box[idx]
    return hpsi


def vloc_psi_gamma(d: Descriptor, lda: int, n: int, m: int, psi: np.ndarray, v: np.ndarray, hpsi: np.ndarray):
    nr = (d.nr1, d.nr2, d.nr3)
    inl = d.nl[:n] - 1
    inlm = d.nlm[:n] - 1
    b = 0
    while b < m:
        box = np.zeros(d.nnr, dtype=np.complex128)
        if b + 1 < m:
            c1, c2 = psi[b, :n], psi[b + 1, :n]
            box[inl] = c1 + 1j * c2
            box[inlm] = np.conj(c1 - 1j * c2)  # written second: wins at G=0
        else:
            c1 = psi[b, :n]
            box[inl] = c1
            box[inlm] = np.conj(c1)
        box = cfft3ds(box, nr, +1, d.isind, d.iplw)
        box = box * v
        box = cfft3ds(box, nr, -1, d.isind, d.iplw)
        if b + 1 < m:
            fp = (box[inl] + box[inlm]) * 0.5
            fm = (box[inl] - box[inlm]) * 0.5
            hpsi[b, :n] += fp.real + 1j * fm.imag
            hpsi[b + 1, :n] += fp.imag - 1j * fm.real
            b += 2
        else:
            hpsi[b, :n] += box[inl]
            b += 1
    return hpsi


def vloc_psi_nc(d: Descriptor, lda: int, n: int, m: int, psi: np.ndarray, v: np.ndarray, hpsi: np.ndarray, domag=True):
    """Noncollinear V_loc, PW/src/vloc_psi_acc.f90:269-369. psi, hpsi: (m, 2*lda) complex = Fortran psi(lda*npol, m) /
    hpsi(lda, npol, m) memory (component ipol of band b at [b, ipol*lda : ipol*lda + n]); v: (4, nnr) = Fortran v(nnr,4)
    memory. hpsi is INOUT. domag: the 2x2 product of :330-340, else v(:,1) on both components (:342-348)."""
    nr = (d.nr1, d.nr2, d.nr3)
    idx = d.nl[d.igk[:n] - 1] - 1
    for b in range(m):
        r = []
        for ipol in range(2):
            box = np.zeros(d.nnr, dtype=np.complex128)
            box[idx] = psi[b, ipol * lda:ipol * lda + n]
            r.append(cfft3ds(box, nr, +1, d.isind, d.iplw))
        if domag:
            sup = r[0] * (v[0] + v[3]) + r[1] * (v[1] - 1j * v[2])
            sdwn = r[1] * (v[0] - v[3]) + r[0] * (v[1] + 1j * v[2])
        else:
            sup, sdwn = r[0] * v[0], r[1] * v[0]
        for ipol, box in enumerate((sup, sdwn)):
            box = cfft3ds(box, nr, -1, d.isind, d.iplw)
            hpsi[b, ipol * lda:ipol * lda + n] += box[idx]
    return hpsi


def apply_deeq_nc(becp_nc: np.ndarray, deeq_nc: np.ndarray, nh_at: np.ndarray, ofsbeta: np.ndarray) -> np.ndarray:
    """ps of add_vuspsi_nc, PW/src/add_vuspsi.f90:192-223. becp_nc (m, 2, nkb) = Fortran becp%nc(nkb,npol,m) memory;
    deeq_nc (4, nat, nhm, nhm) complex = Fortran deeq_nc(nhm,nhm,nat,4) memory, i.e. deeq_nc[is, na, jh, ih]."""
    ps = np.zeros_like(becp_nc)
    for na, (nh, o) in enumerate(zip(nh_at, ofsbeta)):
        D = [deeq_nc[k, na, :nh, :nh].T for k in range(4)]  # D[k][ih, jh]
        b1, b2 = becp_nc[:, 0, o:o + nh], becp_nc[:, 1, o:o + nh]
        ps[:, 0, o:o + nh] += b1 @ D[0].T + b2 @ D[1].T
        ps[:, 1, o:o + nh] += b1 @ D[2].T + b2 @ D[3].T
    return ps


def h_psi_nc(d: Descriptor, lda, n, m, psi, v4, vkb, deeq_nc, nh_at, ofsbeta):
    """Noncollinear h_psi_ with magnetisation (PW/src/h_psi.f90:138-148 kinetic on both components, :184-190
    vloc_psi_nc, :225-229 calbec + add_vuspsi -> calbec_nc becmod.f90:575-616, add_vuspsi_nc add_vuspsi.f90:174-234).
    psi (m, 2*lda); returns (hpsi (m, 2*lda), becp_nc (m, 2, nkb))."""
    hpsi = np.zeros((m, 2 * lda), dtype=np.complex128)
    for ipol in range(2):
        hpsi[:, ipol * lda:ipol * lda + n] = d.g2kin[None, :n] * psi[:, ipol * lda:ipol * lda + n]
    vloc_psi_nc(d, lda, n, m, psi, v4, hpsi)
    nkb = 0 if vkb is None else vkb.shape[0]
    becp = np.zeros((m, 2, nkb), dtype=np.complex128)
    if nkb:
        for ipol in range(2):
            becp[:, ipol, :] = psi[:, ipol * lda:ipol * lda + n] @ np.conj(vkb[:, :n]).T
        ps = apply_deeq_nc(becp, deeq_nc, nh_at, ofsbeta)
        for ipol in range(2):
            hpsi[:, ipol * lda:ipol * lda + n] += ps[:, ipol, :] @ vkb[:, :n]
    return hpsi, becp


def s_psi_nc(lda, n, m, psi, becp_nc, vkb, qq_nc, nh_at, ofsbeta):
    """s_psi_nc, PW/src/s_psi.f90:308-370 (with the copy of all lda*npol rows, :118). qq_nc: (4, nat, nhm, nhm) per atom
    (qq_so of the atom's type, or qq_at in blocks 1 and 4)."""
    spsi = psi.copy()
    if vkb is not None and vkb.shape[0] and qq_nc is not None:
        ps = apply_deeq_nc(becp_nc, qq_nc, nh_at, ofsbeta)
        for ipol in range(2):
            spsi[:, ipol * lda:ipol * lda + n] += ps[:, ipol, :] @ vkb[:, :n]
    return spsi


def synthetic_deeq_nc(nh_at, seed: int = 0):
    """Hermitian 2x2-spin D matrices in the deeq_nc(nhm,nhm,nat,4) layout (as newd_nc / spin-orbit produce them:
    D^{ss'}_{ij} = conj(D^{s's}_{ji})); returned as (4, nat, nhm, nhm) complex."""
    rng = np.random.default_rng(4411 + seed)
    nat, nhm = len(nh_at), int(max(nh_at))
    out = np.zeros((4, nat, nhm, nhm), dtype=np.complex128)
    for na, nh in enumerate(nh_at):
        big = rng.standard_normal((2 * nh, 2 * nh)) + 1j * rng.standard_normal((2 * nh, 2 * nh))
        big = 0.5 * (big + big.conj().T)  # big[(s,ih),(s',jh)]
        for s in range(2):
            for sp in range(2):
                blk = big[s * nh:(s + 1) * nh, sp * nh:(sp + 1) * nh]  # [ih, jh]
                out[2 * s + sp, na, :nh, :nh] = blk.T  # memory [jh, ih]
    return out


def sum_band_rho(d: Descriptor, lda: int, n: int, m: int, evc: np.ndarray, w: np.ndarray, rho: np.ndarray):
    """Density part of sum_band: rho (nnr, INOUT) += sum_b w[b] |psi_b(r)|^2, w[b] = wg(b,ik)/omega.
    k: wave_g2r + get_rho_k per band (PW/src/sum_band.f90:578-593, :650-671); Gamma: two bands per transform,
    rho += w1 Re^2 + w2 Im^2, w2 = w1 for the odd last band (:376-394, :674-696)."""
    nr = (d.nr1, d.nr2, d.nr3)
    if not d.gamma_only:
        idx = d.nl[d.igk[:n] - 1] - 1
        for b in range(m):
            box = np.zeros(d.nnr, dtype=np.complex128)
            box[idx] = evc[b, :n]
            box = cfft3ds(box, nr, +1, d.isind, d.iplw)
            rho += w[b] * (box.real**2 + box.imag**2)
        return rho
    inl, inlm = d.nl[:n] - 1, d.nlm[:n] - 1
    b = 0
    while b < m:
        box = np.zeros(d.nnr, dtype=np.complex128)
        if b + 1 < m:
            c1, c2 = evc[b, :n], evc[b + 1, :n]
            box[inl] = c1 + 1j * c2
            box[inlm] = np.conj(c1 - 1j * c2)
            w1, w2 = w[b], w[b + 1]
        else:
            c1 = evc[b, :n]
            box[inl] = c1
            box[inlm] = np.conj(c1)
            w1 = w2 = w[b]
        box = cfft3ds(box, nr, +1, d.isind, d.iplw)
        rho += w1 * box.real**2 + w2 * box.imag**2
        b += 2
    return rho


def sum_band_rho_nc(d: Descriptor, lda: int, n: int, m: int, evc: np.ndarray, w: np.ndarray, rho: np.ndarray,
                    domag: bool = True):
    """`noncolin` branch of the sum_band band loop (PW/src/sum_band.f90:555-575): one wave_g2r per spinor component
    (evc(1:npw,ibnd), evc(npwx+1:npwx+npw,ibnd)), get_rho_k on both into rho(:,1) (:650-671) and, with domag,
    get_rho_domag into rho(:,2:4) (:699-729). evc: (m, 2*lda); rho: (4, nnr) INOUT, or (nnr,) when not domag."""
    nr = (d.nr1, d.nr2, d.nr3)
    idx = d.nl[d.igk[:n] - 1] - 1
    r1 = rho[0] if domag else rho
    for b in range(m):
        psic = []
        for ipol in range(2):
            box = np.zeros(d.nnr, dtype=np.complex128)
            box[idx] = evc[b, ipol * lda:ipol * lda + n]
            psic.append(cfft3ds(box, nr, +1, d.isind, d.iplw))
        p1, p2 = psic
        for p in psic:
            r1 += w[b] * (p.real**2 + p.imag**2)
        if domag:
            rho[1] += w[b] * 2.0 * (p1.real * p2.real + p1.imag * p2.imag)
            rho[2] += w[b] * 2.0 * (p1.real * p2.imag - p2.real * p1.imag)
            rho[3] += w[b] * (p1.real**2 + p1.imag**2 - p2.real**2 - p2.imag**2)
    return rho


# ----------------------------------------------------------------------------
# Nonlocal
# ----------------------------------------------------------------------------
def calbec_k(n: int, vkb: np.ndarray, psi: np.ndarray, m: int) -> np.ndarray:
    """vkb: (nkb, npwx) rows = projectors; psi: (m, lda). Returns becp (m, nkb) complex
    with becp[j,i] = sum_g conj(vkb[i,g]) psi[j,g]   (becmod.f90:475)."""
    return psi[:m, :n] @ np.conj(vkb[:, :n]).T


def calbec_gamma(n: int, vkb: np.ndarray, psi: np.ndarray, m: int, gstart: int) -> np.ndarray:
    """becp_r[j,i] = 2*sum_g (Re b Re p + Im b Im p) - [gstart==2] Re b(1) Re p(1) (becmod.f90:315-324)."""
    a = vkb[:, :n]
    p = psi[:m, :n]
    r = 2.0 * (p.real @ a.real.T + p.imag @ a.imag.T)
    if gstart == 2:
        r -= np.outer(p[:, 0].real, a[:, 0].real)
    return r


def apply_deeq(becp: np.ndarray, deeq: np.ndarray, nh_at: np.ndarray, ofsbeta: np.ndarray) -> np.ndarray:
    """ps[:, o:o+nh] = becp[:, o:o+nh] @ D_at^T  (add_vuspsi.f90:151-155, :92-97).
    deeq: (nat, nhm, nhm) with deeq[a, jh, ih] = Fortran deeq(ih,jh,a) (i.e. numpy
    array laid out exactly like the Fortran memory)."""
    ps = np.zeros_like(becp)
    for a in range(len(nh_at)):
        nh = int(nh_at[a])
        if nh == 0:
            continue
        o = int(ofsbeta[a])
        D = deeq[a, :nh, :nh].T  # D[ih,jh] = deeq(ih,jh,a)
        ps[:, o:o + nh] = becp[:, o:o + nh] @ D.T
    return ps


def add_vuspsi_k(n, m, vkb, ps, hpsi):
    """hpsi(1:n,:) += vkb(1:n,:) ps  (add_vuspsi.f90:164)."""
    hpsi[:m, :n] += ps[:m] @ vkb[:, :n]
    return hpsi


def add_vuspsi_gamma(n, m, vkb, ps_r, hpsi):
    """real DGEMM on the (2*lda) real view == complex vkb times real ps (add_vuspsi.f90:101-116)."""
    hpsi[:m, :n] += ps_r[:m] @ vkb[:, :n]
    return hpsi


# ----------------------------------------------------------------------------
# h_psi
# ----------------------------------------------------------------------------
def h_psi(d: Descriptor, lda, n, m, psi, vrs, vkb=None, deeq=None, nh_at=None, ofsbeta=None):
    """Returns (hpsi (m,lda), becp (m,nkb) or None)."""
    hpsi = np.zeros((m, lda), dtype=np.complex128)
    hpsi[:, :n] = d.g2kin[None, :n] * psi[:m, :n]
    if d.gamma_only:
        vloc_psi_gamma(d, lda, n, m, psi, vrs, hpsi)
    else:
        vloc_psi_k(d, lda, n, m, psi, vrs, hpsi)
    becp = None
    if vkb is not None and vkb.shape[0] > 0:
        if d.gamma_only:
            becp = calbec_gamma(n, vkb, psi, m, d.gstart)
            ps = apply_deeq(becp, deeq, nh_at, ofsbeta)
            add_vuspsi_gamma(n, m, vkb, ps, hpsi)
        else:
            becp = calbec_k(n, vkb, psi, m)
            ps = apply_deeq(becp, deeq, nh_at, ofsbeta)
            add_vuspsi_k(n, m, vkb, ps, hpsi)
    if d.gamma_only and d.gstart == 2:
        hpsi[:, 0] = hpsi[:, 0].real
    return hpsi, becp


def s_psi(d: Descriptor, lda, n, m, psi, becp, vkb=None, qq_at=None, nh_at=None, ofsbeta=None):
    """s_psi_ (PW/src/s_psi.f90:79-262): spsi = psi on ALL lda rows (threaded_memcpy, :118); with an ultrasoft S,
    ps = blockdiag(qq_at) becp per atom (s_psi_gamma :205-219 DGEMM, s_psi_k :262-276 ZGEMM with qq copied to
    complex) and spsi(1:n) += vkb ps (:222-228, :293-302). becp (m,nkb) is the one h_psi/calbec left (:15)."""
    spsi = np.array(psi[:m], dtype=np.complex128, copy=True)
    if vkb is None or vkb.shape[0] == 0 or qq_at is None:
        return spsi
    ps = apply_deeq(becp, qq_at, nh_at, ofsbeta)
    if d.gamma_only:
        spsi[:m, :n] += ps[:m] @ vkb[:, :n]
    else:
        spsi[:m, :n] += ps[:m] @ vkb[:, :n]
    return spsi


def synthetic_qq(nh_at, seed: int = 0, symmetric: bool = True):
    """qq_at(nhm,nhm,nat) as numpy (nat,nhm,nhm): symmetric, like the integrals of Q_ij (upflib/init_us_1.f90);
    symmetric=False pins the (ih,jh) orientation in a parity test."""
    nh_at = np.asarray(nh_at, dtype=np.int32)
    nhm = int(nh_at.max()) if len(nh_at) else 0
    rng = np.random.default_rng(5234 + seed)
    qq = np.zeros((len(nh_at), nhm, nhm))
    for a, nh in enumerate(nh_at):
        Q = rng.uniform(-0.5, 0.5, (nh, nh))
        qq[a, :nh, :nh] = 0.5 * (Q + Q.T) if symmetric else Q
    return qq


def divide_all(m: int, nproc: int):
    """UtilXlib/divide.f90:55-106 -> (recv_counts, displs) 0-based displs."""
    base, rest = divmod(m, nproc)
    counts = np.array([base + 1 if p < rest else base for p in range(nproc)], dtype=np.int32)
    displs = np.concatenate([[0], np.cumsum(counts)[:-1]]).astype(np.int32)
    return counts, displs


# ----------------------------------------------------------------------------
# Synthetic inputs
# ----------------------------------------------------------------------------
def synthetic_inputs(d: Descriptor, m: int, nh_at, seed: int = 0, lda: int | None = None, poison: bool = True,
                     symmetric: bool = True):
    """Returns dict(psi (m,lda), vrs (nnr), vkb (nkb,lda), deeq (nat,nhm,nhm), nh_at, ofsbeta)."""
    n = d.npw
    if lda is None:
        lda = n + 3
    rng = np.random.default_rng(1234 + seed)
    psi = (rng.standard_normal((m, lda)) + 1j * rng.standard_normal((m, lda))) / np.sqrt(2.0)
    if d.gamma_only and d.gstart == 2:
        psi[:, 0] = psi[:, 0].real
    psi[:, :n] /= np.linalg.norm(psi[:, :n], axis=1, keepdims=True)
    if poison:
        psi[:, n:] = np.nan
    rng = np.random.default_rng(2234 + seed)
    vrs = rng.uniform(-1.0, 0.0, d.nnr)
    nh_at = np.asarray(nh_at, dtype=np.int32)
    nkb = int(nh_at.sum())
    ofsbeta = np.concatenate([[0], np.cumsum(nh_at)[:-1]]).astype(np.int32) if len(nh_at) else np.zeros(0, np.int32)
    rng = np.random.default_rng(3234 + seed)
    vkb = (rng.standard_normal((nkb, lda)) + 1j * rng.standard_normal((nkb, lda))) / np.sqrt(max(n, 1))
    if d.gamma_only and d.gstart == 2 and nkb:
        vkb[:, 0] = vkb[:, 0].real
    if poison and nkb:
        vkb[:, n:] = np.nan
    nhm = int(nh_at.max()) if len(nh_at) else 0
    rng = np.random.default_rng(4234 + seed)
    deeq = np.zeros((len(nh_at), nhm, nhm))
    for a, nh in enumerate(nh_at):
        D = rng.uniform(-1.0, 1.0, (nh, nh))
        # symmetric=False: a general matrix, so that the (ih,jh) orientation of deeq(ih,jh,na) is pinned by the test
        deeq[a, :nh, :nh] = 0.5 * (D + D.T) if symmetric else D
    return dict(psi=psi, vrs=vrs, vkb=vkb, deeq=deeq, nh_at=nh_at, ofsbeta=ofsbeta, lda=lda, n=n, m=m, nkb=nkb)


# ----------------------------------------------------------------------------
# The reference's own C FFT engine (oracle/_ref/libqe_ref.so), when built
# ----------------------------------------------------------------------------
_REF = None


def ref_lib():
    """ctypes handle of oracle/_ref/libqe_ref.so or None."""
    global _REF
    if _REF is None:
        path = os.path.join(_HERE, "_ref", "libqe_ref.so")
        if not os.path.exists(path):
            return None
        L = ctypes.CDLL(path)
        P = ctypes.c_void_p
        I = ctypes.c_int
        L.qeref_cfft3ds.argtypes = [P, I, I, I, I, I, I, I, I, P, P]
        L.qeref_cfft3ds.restype = I
        L.qeref_vloc_psi_k.argtypes = [I, I, I, P, P, P, I, I, I, P, P, P, P, I]
        L.qeref_vloc_psi_k.restype = I
        L.qeref_vloc_psi_gamma.argtypes = [I, I, I, P, P, P, I, I, I, P, P, P, P, I]
        L.qeref_vloc_psi_gamma.restype = I
        _REF = L
    return _REF


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def ref_cfft3ds(f, nr, isign, do_fft_z, do_fft_y, howmany=1):
    L = ref_lib()
    f = np.ascontiguousarray(f, dtype=np.complex128).copy()
    z = np.ascontiguousarray(do_fft_z, dtype=np.int32)
    y = np.ascontiguousarray(do_fft_y, dtype=np.int32)
    rc = L.qeref_cfft3ds(_p(f), nr[0], nr[1], nr[2], nr[0], nr[1], nr[2], howmany, isign, _p(z), _p(y))
    assert rc == 0
    return f


def ref_vloc_psi(d: Descriptor, lda, n, m, psi, v, hpsi, nthreads=0):
    """hpsi (m,lda) INOUT, through the reference FFT engine."""
    L = ref_lib()
    psi = np.ascontiguousarray(psi, dtype=np.complex128)
    v = np.ascontiguousarray(v, dtype=np.float64)
    z = np.ascontiguousarray(d.isind, dtype=np.int32)
    y = np.ascontiguousarray(d.iplw, dtype=np.int32)
    nl = np.ascontiguousarray(d.nl, dtype=np.int32)
    assert hpsi.flags.c_contiguous and hpsi.dtype == np.complex128
    if d.gamma_only:
        nlm = np.ascontiguousarray(d.nlm, dtype=np.int32)
        rc = L.qeref_vloc_psi_gamma(lda, n, m, _p(psi), _p(v), _p(hpsi), d.nr1, d.nr2, d.nr3, _p(nl), _p(nlm), _p(z), _p(y), nthreads)
    else:
        igk = np.ascontiguousarray(d.igk, dtype=np.int32)
        rc = L.qeref_vloc_psi_k(lda, n, m, _p(psi), _p(v), _p(hpsi), d.nr1, d.nr2, d.nr3, _p(nl), _p(igk), _p(z), _p(y), nthreads)
    assert rc == 0
    return hpsi
