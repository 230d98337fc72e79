"""GPU parity of qeb_sum_band_rho_nc (the `noncolin` branch of the sum_band band loop, PW/src/sum_band.f90:555-575,
get_rho_domag :699-729) against the oracle restatement oracle/hpsi_ref.py::sum_band_rho_nc, through the C ABI mirror.
The magnetisation components are formed from differences of scalar densities (density_nc.cu), hence the 1e-11 bound
relative to each component's own maximum (north_star tolerance)."""
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


def comp_err(a, b):
    """max over the components (leading index) of max|a - b| / max|b| of that component"""
    a, b = np.atleast_2d(a), np.atleast_2d(b)
    return max(np.abs(a[k] - b[k]).max() / np.abs(b[k]).max() for k in range(a.shape[0]))


@pytest.mark.parametrize("name", ["si2_small", "ortho_small", "si2", "al4"])
@pytest.mark.parametrize("m,batch", [(1, 1), (5, 2), (6, 4)])
def test_sum_band_rho_nc(mods, name, m, batch):
    qe, D, R, torch = mods
    d = D.config_descriptor(name)
    c = d.extra["config"]
    nh = c.get("nh_list") or [c["nh"]] * c["nat"]
    inp = R.synthetic_inputs(d, 2 * m, nh, seed=40 + m, poison=True)
    lda, n = inp["lda"], inp["n"]
    evc = np.ascontiguousarray(inp["psi"]).reshape(m, 2 * lda)  # pad rows n..lda of both components are NaN
    rng = np.random.default_rng(m)
    w = rng.uniform(0.0, 2.0, m)
    rho0 = rng.uniform(-1.0, 1.0, (4, d.nnr))
    ref = R.sum_band_rho_nc(d, lda, n, m, np.nan_to_num(evc), w, rho0.copy())
    h = qe.HPsi(batch=batch)
    h.setup_from(d, inp)
    # host arrays
    rho = rho0.copy()
    h.sum_band_rho_nc(lda, n, m, evc, w, rho)
    assert comp_err(rho, ref) < TOL
    # the integral of the density: sum_b w_b <psi_b|psi_b> nnr over both components
    nrm = (np.abs(np.nan_to_num(evc).reshape(m, 2, lda)[:, :, :n]) ** 2).sum((1, 2))
    assert abs((rho[0] - rho0[0]).sum() / d.nnr - (w * nrm).sum()) < 1e-10
    # device arrays
    evc_t = torch.from_numpy(evc).cuda()
    rho_t = torch.from_numpy(rho0.copy()).cuda()
    h.sum_band_rho_nc(lda, n, m, evc_t, torch.from_numpy(w).cuda(), rho_t)
    torch.cuda.synchronize()
    assert comp_err(rho_t.cpu().numpy(), ref) < TOL
    # without magnetisation: only rho(:,1), and a (nnr) array is enough
    r1 = rho0[0].copy()
    h.sum_band_rho_nc(lda, n, m, evc, w, r1, domag=False)
    assert comp_err(r1, ref[0]) < 1e-12
    r1_t = torch.from_numpy(rho0[0].copy()).cuda()
    h.sum_band_rho_nc(lda, n, m, evc_t, w, r1_t, domag=False)
    torch.cuda.synchronize()
    assert comp_err(r1_t.cpu().numpy(), ref[0]) < 1e-12
    h.close()


def test_sum_band_rho_nc_edges(mods):
    qe, D, R, torch = mods
    d = D.config_descriptor("si2_small")
    inp = R.synthetic_inputs(d, 4, [4, 4], seed=2, poison=False)
    lda, n = inp["lda"], inp["n"]
    evc = np.ascontiguousarray(inp["psi"]).reshape(2, 2 * lda)
    h = qe.HPsi()
    h.setup_from(d, inp)
    rho = np.ones((4, d.nnr))
    h.sum_band_rho_nc(lda, n, 0, evc, np.zeros(1), rho)  # m = 0: the loop does not execute
    assert np.all(rho == 1.0)
    # a spinor with an empty down component: m_x = m_y = 0, m_z = rho (to rounding of the differences)
    up_only = evc.copy()
    up_only[:, lda:] = 0.0
    w = np.array([1.0, 0.5])
    rho = np.zeros((4, d.nnr))
    h.sum_band_rho_nc(lda, n, 2, up_only, w, rho)
    scale = np.abs(rho[0]).max()
    assert np.abs(rho[1]).max() < 1e-13 * scale and np.abs(rho[2]).max() < 1e-13 * scale
    assert np.abs(rho[3] - rho[0]).max() < 1e-13 * scale
    h.close()
    dg = D.config_descriptor("si2_small_gamma")
    ig = R.synthetic_inputs(dg, 4, [4, 4], seed=2, poison=False)
    hg = qe.HPsi()
    hg.setup_from(dg, ig)
    with pytest.raises(qe.QebError):
        hg.sum_band_rho_nc(ig["lda"], ig["n"], 2, np.ascontiguousarray(ig["psi"]).reshape(2, -1), w, np.zeros((4, dg.nnr)))
    hg.close()
