"""qeb_calbec_static (include/qeb200.h): the allocation-free variant of qeb_calbec, whose split-K scratch is the library's
file-scope device array. Called through the C ABI (ctypes) on device arrays and compared with the existing entry point
qeb_calbec and with the CPU oracle, on two calls in a row (the array is shared and keeps the last call's contents).
Named test_zz_* so that pytest collects it last."""
import ctypes

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
    a2 = np.asarray(a).reshape(np.asarray(a).shape[0], -1)
    b2 = np.asarray(b).reshape(np.asarray(b).shape[0], -1)
    num = np.abs(a2 - b2).max(axis=1)
    den = np.abs(b2).max(axis=1)
    with np.errstate(divide="ignore", invalid="ignore"):
        r = np.where(den > 0, num / den, np.where(num > 0, np.inf, 0.0))
    return float(r.max()) if r.size else 0.0


def _dev(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _call(L, h, fn, n, beta_t, nkb, psi_t, lda, m, out_t):
    P = ctypes.c_void_p
    return getattr(L, fn)(h._h, n, P(beta_t.data_ptr()) if beta_t is not None else None, nkb, P(psi_t.data_ptr()), lda, m,
                          P(out_t.data_ptr()))


@pytest.mark.parametrize("name,nh", [("ortho_small", [8, 8, 8]), ("ortho_small_gamma", [8, 8, 8]), ("si2", [18, 18])])
def test_calbec_static_matches_calbec_twice(mods, name, nh):
    qe, D, R, torch = mods
    d = D.config_descriptor(name)
    h = qe.HPsi()
    L = h._L
    assert L.qeb_static_scratch_bytes() == 32 << 20
    dt = torch.float64 if d.gamma_only else torch.complex128
    # two calls in a row with different band counts and data: the second must not see the first one's partial sums
    for m, seed in ((6, 11), (3, 12), (6, 13)):
        inp = R.synthetic_inputs(d, m, nh, seed=seed, poison=False)
        lda, n, nkb = inp["lda"], inp["n"], inp["nkb"]
        if seed == 11:
            h.setup_from(d, inp)
        vkb_t, psi_t = _dev(torch, inp["vkb"]), _dev(torch, inp["psi"])
        ref = (R.calbec_gamma(n, inp["vkb"], inp["psi"], m, d.gstart) if d.gamma_only
               else R.calbec_k(n, inp["vkb"], inp["psi"], m))
        out_old = torch.zeros((m, nkb), dtype=dt, device="cuda")
        out_new = torch.full((m, nkb), float("nan"), dtype=dt, device="cuda")
        torch.cuda.synchronize()
        assert _call(L, h, "qeb_calbec", n, vkb_t, nkb, psi_t, lda, m, out_old) == 0, h._L.qeb_last_error(h._h).decode()
        rc = _call(L, h, "qeb_calbec_static", n, vkb_t, nkb, psi_t, lda, m, out_new)
        assert rc == 0, h._L.qeb_last_error(h._h).decode()
        torch.cuda.synchronize()
        a, b = out_new.cpu().numpy(), out_old.cpu().numpy()
        assert relerr(a, b) < TOL
        assert relerr(a, ref) < TOL
    # the context's own vkb (beta == NULL), through the Python mirror
    out_ctx = torch.full((m, nkb), float("nan"), dtype=dt, device="cuda")
    inp0 = R.synthetic_inputs(d, m, nh, seed=11, poison=False)
    psi0 = _dev(torch, inp0["psi"])
    ref0 = (R.calbec_gamma(n, inp0["vkb"], inp0["psi"], m, d.gstart) if d.gamma_only
            else R.calbec_k(n, inp0["vkb"], inp0["psi"], m))
    h.calbec_static(n, None, psi0, out_ctx, m, lda=lda)
    torch.cuda.synchronize()
    assert relerr(out_ctx.cpu().numpy(), ref0) < TOL
    h.close()


def test_calbec_static_rejects(mods):
    qe, D, R, torch = mods
    d = D.config_descriptor("ortho_small")
    inp = R.synthetic_inputs(d, 2, [8, 8, 8], seed=3, poison=False)
    lda, n, nkb = inp["lda"], inp["n"], inp["nkb"]
    h = qe.HPsi()
    h.setup_from(d, inp)
    L = h._L
    psi_t = _dev(torch, inp["psi"])
    out_t = torch.zeros((2, nkb), dtype=torch.complex128, device="cuda")
    # host arrays are not staged: that would allocate
    host = np.ascontiguousarray(inp["psi"])
    rc = L.qeb_calbec_static(h._h, n, None, nkb, ctypes.c_void_p(host.ctypes.data), lda, 2, ctypes.c_void_p(out_t.data_ptr()))
    assert rc == 93 and "device pointers" in h._L.qeb_last_error(h._h).decode()
    # beyond the static maximum: nkb*m*16 bytes > 32 MiB even with a single split
    big_nkb, big_m = 2100, 1000
    beta_t = torch.zeros((big_nkb, lda), dtype=torch.complex128, device="cuda")
    psi_b = torch.zeros((big_m, lda), dtype=torch.complex128, device="cuda")
    out_b = torch.zeros((big_m, big_nkb), dtype=torch.complex128, device="cuda")
    torch.cuda.synchronize()
    rc = _call(L, h, "qeb_calbec_static", n, beta_t, big_nkb, psi_b, lda, big_m, out_b)
    assert rc == 75 and "static maximum" in h._L.qeb_last_error(h._h).decode()
    # the context still works afterwards
    assert _call(L, h, "qeb_calbec_static", n, None, nkb, psi_t, lda, 2, out_t) == 0, h._L.qeb_last_error(h._h).decode()
    torch.cuda.synchronize()
    assert relerr(out_t.cpu().numpy(), R.calbec_k(n, inp["vkb"], inp["psi"], 2)) < TOL
    h.close()
