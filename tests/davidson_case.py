"""Shared problem builder of the Davidson tests (CPU oracle test and GPU parity test): a synthetic Kohn-Sham-like
generalized eigenproblem on one of the small descriptors, with the oracle's H and S as the operators."""
import numpy as np

from oracle import davidson_ref as DR
from oracle import descriptor as D
from oracle import hpsi_ref as R


def build(name, nvec, seed=11, nh=(4, 4)):
    d = D.config_descriptor(name)
    inp = R.synthetic_inputs(d, nvec, list(nh), seed=seed, poison=False)
    lda, n = inp["lda"], inp["n"]
    rng = np.random.default_rng(100 + seed)
    # ultrasoft overlap with a positive semi-definite qq, so that S = 1 + sum |beta> qq <beta| is positive definite
    nhm = int(np.max(inp["nh_at"]))
    qq = np.zeros((len(inp["nh_at"]), nhm, nhm))
    for a, k in enumerate(inp["nh_at"]):
        A = rng.uniform(-1, 1, (k, k))
        qq[a, :k, :k] = 0.2 * (A @ A.T)
    vkb, deeq, nh_at, ofs = inp["vkb"], inp["deeq"], inp["nh_at"], inp["ofsbeta"]
    vrs = inp["vrs"]

    def h_op(p):
        out, _ = R.h_psi(d, lda, n, p.shape[0], p, vrs, vkb, deeq, nh_at, ofs)
        return out

    def s_op(p):
        _, becp = R.h_psi(d, lda, n, p.shape[0], p, vrs, vkb, deeq, nh_at, ofs)
        return R.s_psi(d, lda, n, p.shape[0], p, becp, vkb, qq, nh_at, ofs)

    # g_psi_mod diagonals (PW/src/usnldiag.f90: h_diag = g2kin + v_of_0 + sum deeq |vkb|^2, s_diag = 1 + sum qq |vkb|^2)
    h_diag = np.asarray(d.g2kin[:n], dtype=float) + vrs.mean()
    s_diag = np.ones(n)
    for a, k in enumerate(nh_at):
        for ih in range(k):
            b = np.abs(vkb[ofs[a] + ih, :n]) ** 2
            h_diag += deeq[a, ih, ih] * b
            s_diag += qq[a, ih, ih] * b

    def g_op(p, e):
        DR.g_psi(p, e, n, h_diag, s_diag)

    evc = inp["psi"].copy()
    evc[:, n:] = 0.0
    return dict(d=d, inp=inp, lda=lda, n=n, qq=qq, h_op=h_op, s_op=s_op, g_op=g_op, h_diag=h_diag, s_diag=s_diag, evc=evc)


def dense_eigenvalues(case, nvec):
    """lowest eigenvalues of the same (H, S) from dense matrices (Gamma: in the real representation in which the
    reference's 2*dot - G=0 inner product is the Euclidean one)"""
    import scipy.linalg as sl

    d, n, lda = case["d"], case["n"], case["lda"]
    if not d.gamma_only:
        eye = np.zeros((n, lda), complex)
        eye[np.arange(n), np.arange(n)] = 1.0
        H = case["h_op"](eye)[:, :n].T  # column j = H e_j
        S = case["s_op"](eye)[:, :n].T
        H, S = 0.5 * (H + H.conj().T), 0.5 * (S + S.conj().T)
        return sl.eigh(H, S, eigvals_only=True)[:nvec]
    # real coordinates x = [c_0 (real, G=0), sqrt2 Re c_1.., sqrt2 Im c_1..] when gstart == 2
    assert d.gstart == 2
    r2 = np.sqrt(2.0)
    m = 2 * n - 1

    def to_c(x):
        c = np.zeros((x.shape[0], lda), complex)
        c[:, 0] = x[:, 0]
        c[:, 1:n] = (x[:, 1:n] + 1j * x[:, n:]) / r2
        return c

    def to_x(c):
        return np.concatenate([c[:, :1].real, r2 * c[:, 1:n].real, r2 * c[:, 1:n].imag], axis=1)

    basis = to_c(np.eye(m))
    H = to_x(case["h_op"](basis)).T
    S = to_x(case["s_op"](basis)).T
    H, S = 0.5 * (H + H.T), 0.5 * (S + S.T)
    return sl.eigh(H, S, eigvals_only=True)[:nvec]


def build_nc(name, nvec, seed=21, nh=(4, 4)):
    """the same for noncollinear spinors (npol = 2): H = h_psi_nc with a magnetic 2x2 potential and a Hermitian
    deeq_nc, S = s_psi_nc with qq_at on the spin diagonal; vectors are (2*lda) long, pad rows zero"""
    d = D.config_descriptor(name)
    inp = R.synthetic_inputs(d, 2 * nvec, list(nh), seed=seed, poison=False)
    lda, n = inp["lda"], inp["n"]
    rng = np.random.default_rng(200 + seed)
    nh_at, ofs, vkb = inp["nh_at"], inp["ofsbeta"], inp["vkb"]
    nhm = int(np.max(nh_at))
    v4 = np.ascontiguousarray(rng.uniform(-1.0, 1.0, (4, d.nnr)))
    v4[0] = inp["vrs"]
    v4[1:] *= 0.3
    dnc = R.synthetic_deeq_nc(nh_at, seed=seed)
    qq = np.zeros((len(nh_at), nhm, nhm))
    for a, k in enumerate(nh_at):
        A = rng.uniform(-1, 1, (k, k))
        qq[a, :k, :k] = 0.2 * (A @ A.T)
    qnc = np.zeros((4, len(nh_at), nhm, nhm), dtype=np.complex128)
    qnc[0] = qq
    qnc[3] = qq

    def h_op(p):
        out, _ = R.h_psi_nc(d, lda, n, p.shape[0], p, v4, vkb, dnc, nh_at, ofs)
        return out

    def s_op(p):
        _, becp = R.h_psi_nc(d, lda, n, p.shape[0], p, v4, vkb, dnc, nh_at, ofs)
        return R.s_psi_nc(lda, n, p.shape[0], p, becp, vkb, qnc, nh_at, ofs)

    # h_diag(npwx, npol), s_diag(npwx, npol) (usnldiag.f90 with the spin-diagonal blocks of deeq_nc)
    h_diag = np.zeros((2, lda))
    s_diag = np.ones((2, lda))
    for ipol in range(2):
        h_diag[ipol, :n] = np.asarray(d.g2kin[:n], dtype=float) + v4[0].mean()
        for a, k in enumerate(nh_at):
            for ih in range(k):
                b = np.abs(vkb[ofs[a] + ih, :n]) ** 2
                h_diag[ipol, :n] += dnc[3 * ipol, a, ih, ih].real * b
                s_diag[ipol, :n] += qq[a, ih, ih] * b

    def g_op(p, e):
        DR.g_psi_nc(p, e, n, lda, h_diag, s_diag)

    evc = np.ascontiguousarray(inp["psi"]).reshape(nvec, 2 * lda).copy()
    evc[:, n:lda] = 0.0
    evc[:, lda + n:] = 0.0
    return dict(d=d, inp=inp, lda=lda, n=n, v4=v4, dnc=dnc, qnc=qnc, h_op=h_op, s_op=s_op, g_op=g_op, h_diag=h_diag,
                s_diag=s_diag, evc=evc)


def dense_eigenvalues_nc(case, nvec):
    import scipy.linalg as sl

    n, lda = case["n"], case["lda"]
    rows = np.concatenate([np.arange(n), lda + np.arange(n)])
    eye = np.zeros((2 * n, 2 * lda), complex)
    eye[np.arange(2 * n), rows] = 1.0
    H = case["h_op"](eye)[:, rows].T
    S = case["s_op"](eye)[:, rows].T
    assert np.abs(H - H.conj().T).max() < 1e-10 * np.abs(H).max()
    assert np.abs(S - S.conj().T).max() < 1e-10 * np.abs(S).max()
    H, S = 0.5 * (H + H.conj().T), 0.5 * (S + S.conj().T)
    return sl.eigh(H, S, eigvals_only=True)[:nvec]
