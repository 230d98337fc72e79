"""CPU tests of the N>1 host logic (no GPU): the restated R&G distribution (oracle/dist.py), its data flow against
the serial oracle, and - with two real processes over gloo - the stick<->plane exchange driven only by rank-local
descriptor data plus the band-group all-gather of the Python mirror (qe_b200.allgather_columns)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import descriptor as D  # noqa: E402
from oracle import dist as X  # noqa: E402
from oracle import hpsi_ref as R  # noqa: E402

CASES = ["ortho_small", "ortho_small_gamma", "si2_small", "si2_small_gamma"]


@pytest.mark.parametrize("name", CASES)
@pytest.mark.parametrize("nproc", [1, 2, 3, 4, 8])
def test_distribution_properties(name, nproc):
    d = D.config_descriptor(name)
    if d.nr3 < nproc:
        pytest.skip("fewer planes than ranks")
    dd = X.distribute(d, nproc)
    assert dd.nr3p.sum() == d.nr3 and dd.nr3p.max() - dd.nr3p.min() <= 1
    assert dd.nsw.sum() == d.nst                      # every wave stick of the serial descriptor, owned once
    wave_cols = np.concatenate([dd.ismap[dd.iss[p]:dd.iss[p] + dd.nsw[p]] for p in range(nproc)]) - 1
    assert len(set(wave_cols.tolist())) == d.nst
    assert np.array_equal(np.sort(wave_cols), np.nonzero(d.isind)[0])
    assert sum(rk.npw for rk in dd.ranks) == d.npw    # every plane wave on exactly one rank
    assert sum(rk.ngm for rk in dd.ranks) == d.ngm
    rows = np.concatenate([rk.pw_l2g for rk in dd.ranks])
    assert np.array_equal(np.sort(rows), np.arange(d.npw))
    assert sum(1 for rk in dd.ranks if rk.gstart == 2) == (1 if d.gstart == 2 else 0)
    if d.gamma_only:                                  # +G and -G sticks live on the same rank
        own = np.abs(dd.isind_owner).reshape(d.nr2, d.nr1)
        for y in range(d.nr2):
            for x in range(d.nr1):
                if own[y, x]:
                    assert own[(-y) % d.nr2, (-x) % d.nr1] == own[y, x]
    # the greedy rule balances G-vectors: no rank is more than one longest stick away from the mean
    npw = np.array([rk.npw for rk in dd.ranks])
    assert npw.max() - npw.min() <= 2 * d.nr3


@pytest.mark.parametrize("name", CASES)
@pytest.mark.parametrize("nproc", [2, 3])
def test_distributed_flow_equals_serial(name, nproc):
    d = D.config_descriptor(name)
    m = 5
    inp = R.synthetic_inputs(d, m, [8, 8, 8], seed=5, poison=False)
    n = inp["n"]
    ref, becp_ref = R.h_psi(d, inp["lda"], n, m, inp["psi"], inp["vrs"], inp["vkb"], inp["deeq"], inp["nh_at"], inp["ofsbeta"])
    dd = X.distribute(d, nproc)
    h, becp = X.dist_h_psi(dd, m, X.split_rows(dd, inp["psi"][:, :n]), X.v_slabs(dd, inp["vrs"]),
                           X.split_rows(dd, inp["vkb"][:, :n]), inp["deeq"], inp["nh_at"], inp["ofsbeta"])
    for r, rk in enumerate(dd.ranks):
        assert np.abs(h[r][:, :rk.npw] - ref[:, rk.pw_l2g]).max() / np.abs(ref).max() < 1e-13
    assert np.abs(becp - becp_ref).max() / np.abs(becp_ref).max() < 1e-13


def _gloo_worker(rank, world, port, name, resq):
    import torch
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import qe_b200

        d = D.config_descriptor(name)
        m = 5
        inp = R.synthetic_inputs(d, m, [8, 8, 8], seed=5, poison=False)
        n = inp["n"]
        ref, _ = R.h_psi(d, inp["lda"], n, m, inp["psi"], inp["vrs"])
        dd = X.distribute(d, world)
        rk = dd.ranks[rank]
        # (1) V_loc with a real exchange: this process only touches its own rows/sticks/planes
        psi_l = X.split_rows(dd, inp["psi"][:, :n])[rank]
        v_l = X.v_slabs(dd, inp["vrs"])[rank]
        hp = np.zeros_like(psi_l)
        hp[:, :rk.npw] = rk.g2kin[None, :] * psi_l[:, :rk.npw]
        nr1, nr2, nr3 = d.nr1, d.nr2, d.nr3
        mycols = dd.ismap[dd.iss[rank]:dd.iss[rank] + dd.nsw[rank]] - 1
        for b in range(m):
            f = np.zeros(int(dd.nsw[rank]) * nr3, dtype=complex)
            f[rk.nl[rk.igk - 1] - 1] = psi_l[b, :rk.npw]
            s = np.fft.ifft(f.reshape(-1, nr3), axis=1) * nr3
            send = [np.ascontiguousarray(s[:, dd.i0r3p[q]:dd.i0r3p[q] + dd.nr3p[q]]) for q in range(world)]
            recv = [None] * world
            for q in range(world):  # all-to-all through gather (gloo has no all_to_all)
                got = [None] * world
                dist.all_gather_object(got, send[q])
                # got[p] = what p prepared for q in this round; keep the block addressed to me
                if q == rank:
                    recv = got
            pl = np.zeros((int(dd.nr3p[rank]), nr1 * nr2), dtype=complex)
            for p in range(world):
                cols = dd.ismap[dd.iss[p]:dd.iss[p] + dd.nsw[p]] - 1
                pl[:, cols] = recv[p].T
            box = np.fft.ifft2(pl.reshape(-1, nr2, nr1), axes=(1, 2)) * (nr1 * nr2)
            box = np.fft.fft2(box * v_l, axes=(1, 2)) / (nr1 * nr2)
            pl = box.reshape(-1, nr1 * nr2)
            back_send = [np.ascontiguousarray(pl[:, dd.ismap[dd.iss[p]:dd.iss[p] + dd.nsw[p]] - 1].T) for p in range(world)]
            s2 = np.zeros_like(s)
            for p in range(world):
                got = [None] * world
                dist.all_gather_object(got, back_send[p])
                if p == rank:
                    for q in range(world):
                        s2[:, dd.i0r3p[q]:dd.i0r3p[q] + dd.nr3p[q]] = got[q]
            f2 = (np.fft.fft(s2, axis=1) / nr3).reshape(-1)
            hp[b, :rk.npw] += f2[rk.nl[rk.igk - 1] - 1]
        err = np.abs(hp[:, :rk.npw] - ref[:, rk.pw_l2g]).max() / np.abs(ref).max()
        assert mycols.size == dd.nsw[rank]
        # (2) band-group all-gather of the mirror (h_psi.f90:58-63) on CPU tensors
        full = torch.from_numpy(np.arange(7 * 6, dtype=np.float64).reshape(7, 6).copy())
        mine = torch.zeros_like(full)
        s0, e0 = qe_b200.divide_all(7, world, rank)
        mine[s0:e0] = full[s0:e0]
        qe_b200.allgather_columns(mine, 7)
        resq.put((rank, float(err), bool(torch.equal(mine, full))))
    finally:
        dist.destroy_process_group()


def test_two_process_gloo_exchange_and_allgather():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29700 + (os.getpid() % 1000)
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, "ortho_small", q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, err, gathered in res:
        assert err < 1e-13, (rank, err)
        assert gathered


def test_divide_all_matches_reference_rule():
    import qe_b200

    for m in (0, 1, 7, 16, 800):
        for nproc in (1, 2, 3, 4, 8):
            blocks = [qe_b200.divide_all(m, nproc, r) for r in range(nproc)]
            counts, displs = R.divide_all(m, nproc)
            assert [b[0] for b in blocks] == displs.tolist() and [b[1] - b[0] for b in blocks] == counts.tolist()
            assert blocks[0][0] == 0 and blocks[-1][1] == m
            assert all(blocks[i][1] == blocks[i + 1][0] for i in range(nproc - 1))
