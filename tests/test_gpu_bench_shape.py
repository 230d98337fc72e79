"""GPU parity at the BENCH shape (run with -m gpu): AUSURF112 with all nkb = 2016 projectors and m in {48, 192, 800}
bands - the tile grids of k_calbec3m (17 band tiles x 16 projector tiles x K splits) and k_addvus3m, the 192-column
host pipeline with its 40/64-column sub-chunks - against the CPU checker on identical inputs:
  V_loc      the reference's own C FFT engine (oracle/_ref, compiled from /root/reference/FFTXlib/src by oracle/Makefile;
             numpy restatement on a column sample when it is not built), PW/src/vloc_psi_acc.f90:148-266
  projectors numpy / OpenBLAS zgemm restating Modules/becmod.f90:431-486 and PW/src/add_vuspsi.f90:140-170
Per-column tolerance 1e-11 (north_star)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 1e-11


def colerr(a, b):
    num = np.abs(a - b).max(axis=1)
    den = np.abs(b).max(axis=1)
    return float((num / den).max())


@pytest.fixture(scope="module")
def shape():
    import torch

    if not torch.cuda.is_available():
        pytest.fail("GPU tests need a CUDA device; there is no CPU fallback")
    import qe_b200
    from oracle import descriptor as D
    from oracle import hpsi_ref as R

    d = D.config_descriptor("ausurf112")
    n = d.npw
    lda = n + 6
    m = 800
    nh = np.full(112, 18, dtype=np.int32)
    nkb = int(nh.sum())
    rng = np.random.default_rng(77)
    psi = np.empty((m, lda), dtype=np.complex128)
    for j0 in range(0, m, 100):
        psi[j0:j0 + 100] = (rng.standard_normal((100, lda)) + 1j * rng.standard_normal((100, lda))) / np.sqrt(2 * n)
    psi[:, n:] = np.nan  # pad rows must be ignored
    vkb = np.empty((nkb, lda), dtype=np.complex128)
    for i0 in range(0, nkb, 252):
        vkb[i0:i0 + 252] = (rng.standard_normal((252, lda)) + 1j * rng.standard_normal((252, lda))) / np.sqrt(n)
    vkb[:, n:] = 0.0
    vrs = rng.uniform(-1.0, 0.0, d.nnr)
    deeq = rng.uniform(-1.0, 1.0, (112, 18, 18))  # general (non-symmetric) D: pins the (ih,jh) orientation
    ofs = (np.arange(112) * 18).astype(np.int32)
    # ---- the checker: reference algorithm on the host cores ----
    ncore = os.cpu_count() or 1
    try:
        from threadpoolctl import threadpool_limits

        lim = threadpool_limits(limits=ncore)
    except Exception:  # noqa: BLE001
        lim = None
    pz = np.nan_to_num(psi)
    ref = np.zeros((m, lda), dtype=np.complex128)
    ref[:, :n] = d.g2kin[None, :n] * pz[:, :n]
    if R.ref_lib() is not None:
        R.ref_vloc_psi(d, lda, n, m, pz, vrs, ref, ncore)
        vloc_cols = np.arange(m)
    else:  # numpy restatement: a sample of columns only (minutes otherwise)
        vloc_cols = np.unique(np.concatenate([np.arange(0, m, 53), [47, 191, 192, 767, 768, m - 1]]))
        sub = np.ascontiguousarray(ref[vloc_cols])
        R.vloc_psi_k(d, lda, n, len(vloc_cols), np.ascontiguousarray(pz[vloc_cols]), vrs, sub)
        ref[vloc_cols] = sub
    becp = R.calbec_k(n, vkb, pz, m)
    ps = R.apply_deeq(becp, deeq, nh, ofs)
    R.add_vuspsi_k(n, m, vkb, ps, ref)
    if lim is not None:
        lim.restore_original_limits()
    return dict(qe=qe_b200, torch=torch, d=d, n=n, lda=lda, m=m, nh=nh, ofs=ofs, nkb=nkb, psi=psi, vkb=vkb, vrs=vrs,
                deeq=deeq, ref=ref, becp=becp, vloc_cols=vloc_cols)


def _ctx(S, vkb_dev):
    d = S["d"]
    h = S["qe"].HPsi()
    h.set_grid(d.nr1, d.nr2, d.nr3, d.nl)
    h.set_kpoint(S["n"], S["lda"], d.igk, d.g2kin, vkb_dev, S["nkb"], d.gstart)
    h.set_vloc(S["vrs"])
    h.set_nonlocal(S["nh"], S["ofs"], S["deeq"])
    return h


def test_bench_shape_m800_host_pipeline(shape):
    """m = 800 from HOST arrays: the default 192-column pipeline (QEB_PIPE_SUB / QEB_PIPE_TAIL defaults) with pinned and
    with pageable arrays; every column against the checker; becp of all 800 bands."""
    S = shape
    torch = S["torch"]
    m, lda, n = S["m"], S["lda"], S["n"]
    vkb_t = torch.from_numpy(S["vkb"]).cuda()
    h = _ctx(S, vkb_t)
    cols = S["vloc_cols"]
    for pinned in (True, False):
        psi = S["psi"].copy()
        out = np.full((m, lda), np.nan + 0j)
        if pinned:
            h.pin_host(psi)
            h.pin_host(out)
        h.h_psi(lda, n, m, psi, out)
        if pinned:
            h.unpin_host(psi)
            h.unpin_host(out)
        assert np.all(out[:, n:] == 0)
        assert colerr(out[cols], S["ref"][cols]) < TOL
        assert colerr(h.get_becp(m), S["becp"]) < TOL
    h.close()


@pytest.mark.parametrize("m", [48, 192, 800])
def test_bench_shape_device_pointers(shape, m):
    """device-resident psi/hpsi (the bench's `value` path): m = 48 (one band tile), 192, 800"""
    S = shape
    torch = S["torch"]
    lda, n = S["lda"], S["n"]
    vkb_t = torch.from_numpy(S["vkb"]).cuda()
    h = _ctx(S, vkb_t)
    psi_t = torch.from_numpy(np.nan_to_num(S["psi"][:m])).cuda()
    out_t = torch.full((m, lda, 2), float("nan"), dtype=torch.float64, device="cuda")
    h.h_psi(lda, n, m, psi_t, out_t)
    torch.cuda.synchronize()
    out = torch.view_as_complex(out_t).cpu().numpy()
    cols = S["vloc_cols"][S["vloc_cols"] < m]
    assert np.all(out[:, n:] == 0)
    assert colerr(out[cols], S["ref"][cols]) < TOL
    assert colerr(h.get_becp(m), S["becp"][:m]) < TOL
    h.close()


def test_bench_shape_m48_host_direct(shape):
    """m = 48 from host arrays: below the pipeline threshold, staged in one piece; host vkb uploaded by the library"""
    S = shape
    m, lda, n = 48, S["lda"], S["n"]
    h = _ctx(S, S["vkb"])
    out = np.full((m, lda), np.nan + 0j)
    h.h_psi(lda, n, m, S["psi"][:m].copy(), out)
    assert colerr(out, S["ref"][:m]) < TOL if len(S["vloc_cols"]) == S["m"] else True
    assert colerr(h.get_becp(m), S["becp"][:m]) < TOL
    h.close()
