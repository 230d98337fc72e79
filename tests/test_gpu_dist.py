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


@pytest.mark.parametrize("name", ["ortho_small", "ortho_small_gamma", "si2_small", "si2_small_gamma"])
@pytest.mark.parametrize("m,batch", [(5, 0), (4, 3), (1, 0)])
def test_dist_single_rank(mods, name, m, batch):
    qe, D, R, X, torch = mods
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from dist_worker import setup_rank

    d = D.config_descriptor(name)
    inp = R.synthetic_inputs(d, m, [8, 8, 8], seed=13, poison=False)
    n = inp["n"]
    ref, becp_ref = R.h_psi(d, inp["lda"], n, m, inp["psi"], inp["vrs"], inp["vkb"], inp["deeq"], inp["nh_at"], inp["ofsbeta"])
    dd = X.distribute(d, 1)
    rk = dd.ranks[0]
    h = qe.HPsi(batch=batch or None)
    lda = setup_rank(h, d, dd, 0, inp, pad=2)
    psi_l = X.split_rows(dd, inp["psi"][:, :n], pad=2, fill=np.nan)[0]
    out = np.full((m, lda), np.nan + 1j * np.nan)
    h.h_psi(lda, rk.npw, m, psi_l, out)
    assert np.abs(out[:, :rk.npw] - ref[:, rk.pw_l2g]).max() / np.abs(ref).max() < TOL
    assert np.all(out[:, rk.npw:] == 0)
    assert np.abs(h.get_becp(m) - becp_ref).max() / np.abs(becp_ref).max() < TOL
    h.close()


def _run_worker(nproc, name, m):
    port = 29500 + (os.getpid() % 2000)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "dist_worker.py"), name, str(m)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    return r.returncode, r.stdout + r.stderr


@pytest.mark.parametrize("name,m", [("ortho_small", 5), ("ortho_small_gamma", 5), ("si2_small_gamma", 3)])
def test_dist_multi_gpu(mods, name, m):
    qe, D, R, X, torch = mods
    ngpu = torch.cuda.device_count()
    if ngpu < 2:
        pytest.skip("needs >= 2 GPUs (gpurun --gpus 2)")
    rc, out = _run_worker(min(ngpu, 4), name, m)
    assert rc == 0 and "DIST_OK" in out, out[-3000:]
