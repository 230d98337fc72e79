"""GPU parity tests of the slab-distributed (R&G) path and of the NCCL collectives (run with -m gpu).
nproc=1 runs in-process on one GPU (exchange = device copy, pack/unpack and the two table sets are exercised);
nproc=2..N needs that many GPUs and runs tests/dist_worker.py under torch.distributed.run."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 1e-11
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def mods():
    import torch

    if not torch.cuda.is_available():
        pytest.fail("GPU tests need a CUDA device; there is no CPU fallback")
    import qe_b200
    from oracle import descriptor as D
    from oracle import dist as X
    from oracle import hpsi_ref as R

    return qe_b200, D, R, X, torch


def _colerr(a, b):
    num = np.abs(a - b).max(axis=1)
    den = np.abs(b).max(axis=1)
    return float((num / den).max())


@pytest.mark.parametrize("name,m,batch", [(nm, m, b) for nm in ("ortho_small", "ortho_small_gamma", "si2_small", "si2_small_gamma")
                                          for m, b in ((5, 0), (4, 3), (1, 0))] +
                         [("si512", 5, 0), ("h2o64", 4, 0), ("ausurf112", 3, 2)])  # BASELINE grids on the slab path
def test_dist_single_rank(mods, name, m, batch):
    qe, D, R, X, torch = mods
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from dist_worker import setup_rank

    d = D.config_descriptor(name)
    nh = [4] * 6 if name in ("si512", "h2o64", "ausurf112") else [8, 8, 8]
    inp = R.synthetic_inputs(d, m, nh, seed=13, poison=False)
    n = inp["n"]
    ref, becp_ref = R.h_psi(d, inp["lda"], n, m, inp["psi"], inp["vrs"], inp["vkb"], inp["deeq"], inp["nh_at"], inp["ofsbeta"])
    dd = X.distribute(d, 1)
    rk = dd.ranks[0]
    h = qe.HPsi(batch=batch or None)
    lda = setup_rank(h, d, dd, 0, inp, pad=2)
    psi_l = X.split_rows(dd, inp["psi"][:, :n], pad=2, fill=np.nan)[0]
    out = np.full((m, lda), np.nan + 1j * np.nan)
    h.h_psi(lda, rk.npw, m, psi_l, out)
    assert _colerr(out[:, :rk.npw], ref[:, rk.pw_l2g]) < TOL
    assert np.all(out[:, rk.npw:] == 0)
    assert _colerr(h.get_becp(m), becp_ref) < TOL
    # density part of sum_band on the distributed descriptor: the rank's own z planes (here: all of them)
    rng = np.random.default_rng(3)
    w = rng.uniform(0.0, 2.0, m)
    rho0 = rng.uniform(0.0, 1.0, d.nnr)
    rho_ref = R.sum_band_rho(d, inp["lda"], n, m, inp["psi"], w, rho0.copy())
    rho = rho0.copy()
    h.sum_band_rho(lda, rk.npw, m, np.nan_to_num(psi_l), w, rho)
    assert np.abs(rho - rho_ref).max() / np.abs(rho_ref).max() < 1e-12
    h.close()


def _run_worker(nproc, name, m):
    port = 29500 + (os.getpid() % 2000)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "dist_worker.py"), name, str(m)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    out = r.stdout + r.stderr
    try:  # post-mortem copy (gpurun brings gpurun_out/ back)
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", f"dist_worker_{name}_n{nproc}.log"), "w") as f:
            f.write(out)
    except OSError:
        pass
    return r.returncode, out


@pytest.mark.parametrize("name,m", [("ortho_small", 5), ("ortho_small_gamma", 5), ("si2_small_gamma", 3), ("si512", 5),
                                    ("h2o64", 4)])
def test_dist_multi_gpu(mods, name, m):
    """every GPU of the box is one rank of the R&G group (up to 8): NCCL exchange, peer-to-peer exchange, band groups;
    si512 / h2o64 = BASELINE configs 5 and 3 at full grid size (reduced projector count)"""
    qe, D, R, X, torch = mods
    ngpu = torch.cuda.device_count()
    if ngpu < 2:
        pytest.skip("needs >= 2 GPUs (gpurun --gpus 2)")
    nproc = min(ngpu, 8)
    d = D.config_descriptor(name)
    while d.nr3 < nproc:  # every rank needs a plane (fft_types.f90:633-635)
        nproc //= 2
    rc, out = _run_worker(nproc, name, m)
    lines = [ln for ln in out.splitlines() if ("rank" in ln and "err" in ln) or "Error" in ln or "failed" in ln]
    assert rc == 0 and "DIST_OK" in out, "\n".join(lines[-40:]) + "\n" + out[-1500:]
