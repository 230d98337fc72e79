"""GPU parity of the launch-bound path: CUDA-graph replay of repeated h_psi calls and qeb_h_psi_multi (several k-points in
flight, the k loop of PW/src/c_bands.f90:95-150) against the oracle. The eager, capture and replay generations of the same
call must all give the oracle's result, and a state change between calls (new potential, new k-point) must not be served
from a stale graph."""
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


def _colerr(a, b):
    return float((np.abs(a - b).max(axis=1) / np.abs(b).max(axis=1)).max())


@pytest.mark.parametrize("name", ["si2", "si2_small_gamma", "al4"])
def test_graph_replay_and_state_changes(mods, name, monkeypatch):
    qe, D, R, torch = mods
    monkeypatch.setenv("QEB_GRAPH", "2")  # graphs for every device-pointer call
    d = D.config_descriptor(name)
    m = 4
    inp = R.synthetic_inputs(d, m, [4, 4], seed=31, poison=False)
    lda, n = inp["lda"], inp["n"]
    h = qe.HPsi()
    h.setup_from(d, inp)
    psi_t = torch.from_numpy(inp["psi"]).cuda()
    out_t = torch.empty_like(psi_t)

    def ref_for(vrs, deeq):
        return R.h_psi(d, lda, n, m, inp["psi"], vrs, inp["vkb"], deeq, inp["nh_at"], inp["ofsbeta"])

    ref, becp_ref = ref_for(inp["vrs"], inp["deeq"])
    for gen in range(6):  # eager (sizes buffers), eager (final key), capture, replay, replay, replay
        out_t.fill_(float("nan"))
        h.h_psi(lda, n, m, psi_t, out_t)
        torch.cuda.synchronize()
        assert _colerr(out_t.cpu().numpy()[:, :n], ref[:, :n]) < TOL, gen
        assert _colerr(h.get_becp(m), becp_ref) < TOL, gen
    # a new potential and new D_ij: the replayed launches must see them (same buffers) or be re-captured (new epoch)
    vrs2 = np.random.default_rng(5).uniform(-1, 0, d.nnr)
    deeq2 = inp["deeq"] * 0.5
    h.set_vloc(vrs2)
    h.set_nonlocal(inp["nh_at"], inp["ofsbeta"], deeq2)
    ref2, _ = ref_for(vrs2, deeq2)
    for gen in range(4):
        h.h_psi(lda, n, m, psi_t, out_t)
        torch.cuda.synchronize()
        assert _colerr(out_t.cpu().numpy()[:, :n], ref2[:, :n]) < TOL, gen
    # other arguments: fewer bands, another output buffer
    out2_t = torch.full_like(psi_t, float("nan"))
    for gen in range(4):
        h.h_psi(lda, n, m - 1, psi_t, out2_t)
        torch.cuda.synchronize()
        assert _colerr(out2_t.cpu().numpy()[: m - 1, :n], ref2[: m - 1, :n]) < TOL
    assert torch.isnan(out2_t[m - 1]).all()
    h.close()


def test_h_psi_multi_k_points(mods):
    """several k-points of the Si-2 cell in one call, each with its own context (own igk, g2kin, vkb, stream)"""
    qe, D, R, torch = mods
    base = D.CONFIGS["si2"]
    nk, m = 5, 4
    ctxs, psis, outs, refs, ns = [], [], [], [], []
    for k in range(nk):
        xk = np.array(base["xk"], dtype=float) + 0.05 * k
        d = D.build_descriptor(base["at"], base["alat"], base["ecutwfc"], xk=xk, gamma_only=False, ecutrho=base["ecutrho"])
        inp = R.synthetic_inputs(d, m, [4, 4], seed=40 + k, poison=False)
        h = qe.HPsi()
        h.setup_from(d, inp)
        ctxs.append(h)
        psis.append(torch.from_numpy(inp["psi"]).cuda())
        outs.append(torch.empty_like(psis[-1]))
        ref, _ = R.h_psi(d, inp["lda"], inp["n"], m, inp["psi"], inp["vrs"], inp["vkb"], inp["deeq"], inp["nh_at"], inp["ofsbeta"])
        refs.append(ref)
        ns.append((inp["lda"], inp["n"]))
    for gen in range(5):
        for o in outs:
            o.fill_(float("nan"))
        torch.cuda.synchronize()
        qe.h_psi_multi(ctxs, [a for a, _ in ns], [b for _, b in ns], [m] * nk, psis, outs)
        for k in range(nk):
            n = ns[k][1]
            assert _colerr(outs[k].cpu().numpy()[:, :n], refs[k][:, :n]) < TOL, (gen, k)
    for h in ctxs:
        h.close()
