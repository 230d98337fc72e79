"""Worker for the multi-GPU parity tests: one process per GPU (launched by torch.distributed.run), every rank
holds its own rows of psi/vkb and its own z planes of V(r), runs the R&G-distributed h_psi through the C ABI and
compares its rows with the serial CPU oracle. Also checks the band-group path (h_psi_bgrp + NCCL all-gather).
Prints 'DIST_OK' from rank 0 when every rank agrees. usage: dist_worker.py CONFIG M"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import qe_b200  # noqa: E402
from oracle import descriptor as D  # noqa: E402
from oracle import dist as X  # noqa: E402
from oracle import hpsi_ref as R  # noqa: E402

TOL = 1e-11


def colerr(a, b):
    """per-column (band) error relative to that column's own max"""
    num = np.abs(a - b).max(axis=1)
    den = np.abs(b).max(axis=1)
    return float((num / den).max())


def setup_rank(h, d, dd, me, inp, pad):
    rk = dd.ranks[me]
    h.set_grid_dist(d.nr1, d.nr2, d.nr3, dd.nproc, me, dd.nr3p, dd.nsw, dd.iss, dd.ismap, rk.nl,
                    rk.nlm if d.gamma_only else None, d.gamma_only)
    n = inp["n"]
    vkb_l = X.split_rows(dd, inp["vkb"][:, :n], pad=pad)[me]
    lda = rk.npw + pad
    h.set_kpoint(rk.npw, lda, None if d.gamma_only else rk.igk, rk.g2kin, vkb_l, vkb_l.shape[0], rk.gstart)
    h.set_vloc(np.ascontiguousarray(X.v_slabs(dd, inp["vrs"])[me]).reshape(-1))
    h.set_nonlocal(inp["nh_at"], inp["ofsbeta"], inp["deeq"])
    return lda


def main():
    name, m = sys.argv[1], int(sys.argv[2])
    big = name in ("si512", "h2o64", "ausurf112")  # BASELINE grids: fewer projectors keep the CPU oracle in seconds
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    d = D.config_descriptor(name)
    nh = [4] * 6 if big else [8, 8, 8]
    inp = R.synthetic_inputs(d, m, nh, seed=21, poison=False)
    n = inp["n"]
    ref, becp_ref = R.h_psi(d, inp["lda"], n, m, inp["psi"], inp["vrs"], inp["vkb"], inp["deeq"], inp["nh_at"], inp["ofsbeta"])
    ok = True
    # ---- R&G distribution over all ranks ----
    dd = X.distribute(d, world)
    rk = dd.ranks[rank]
    h = qe_b200.HPsi(device=local)
    h.nccl_init(qe_b200.COMM_RG)
    for batch in (0, 2):
        if batch:
            h.set_batch(batch)
        lda = setup_rank(h, d, dd, rank, inp, pad=3)
        psi_l = X.split_rows(dd, inp["psi"][:, :n], pad=3, fill=np.nan)[rank]
        out = np.full((m, lda), np.nan + 1j * np.nan)
        h.h_psi(lda, rk.npw, m, psi_l, out)
        # per band, relative to the band's own norm over ALL plane waves (this rank holds a subset of the rows)
        e1 = float((np.abs(out[:, :rk.npw] - ref[:, rk.pw_l2g]).max(axis=1) / np.abs(ref).max(axis=1)).max())
        e2 = colerr(h.get_becp(m), becp_ref)
        ok = ok and e1 < TOL and e2 < TOL and np.all(out[:, rk.npw:] == 0)
        print(f"rank {rank} R&G batch={batch}: hpsi err {e1:.2e} becp err {e2:.2e}", flush=True)
    h.close()
    # ---- the same path with the peer-to-peer exchange (cudaIpc buffers, copy-engine puts, flag words) ----
    hp = qe_b200.HPsi(device=local, batch=2)  # m > batch: several batches alternate between the two slots
    hp.nccl_init(qe_b200.COMM_RG)             # becp all-reduce
    lda = setup_rank(hp, d, dd, rank, inp, pad=3)
    hp.p2p_setup()
    psi_l = X.split_rows(dd, inp["psi"][:, :n], pad=3, fill=np.nan)[rank]
    for rep in range(3):  # repeated calls: epochs keep counting
        out = np.full((m, lda), np.nan + 1j * np.nan)
        hp.h_psi(lda, rk.npw, m, psi_l, out)
        e1 = float((np.abs(out[:, :rk.npw] - ref[:, rk.pw_l2g]).max(axis=1) / np.abs(ref).max(axis=1)).max())
        ok = ok and e1 < TOL and np.all(out[:, rk.npw:] == 0)
    print(f"rank {rank} R&G p2p exchange: hpsi err {e1:.2e}", flush=True)
    # the unfused sequence (S staging + copy-engine puts + repack kernels) on the same context: on the large grids the
    # default is the fused exchange (z pass stores into the peers' buffers, y passes read / write the exchange buffer)
    os.environ["QEB_FUSED_X"] = "0"
    out = np.full((m, lda), np.nan + 1j * np.nan)
    hp.h_psi(lda, rk.npw, m, psi_l, out)
    e1u = float((np.abs(out[:, :rk.npw] - ref[:, rk.pw_l2g]).max(axis=1) / np.abs(ref).max(axis=1)).max())
    ok = ok and e1u < TOL and np.all(out[:, rk.npw:] == 0)
    del os.environ["QEB_FUSED_X"]
    out = np.full((m, lda), np.nan + 1j * np.nan)
    hp.h_psi(lda, rk.npw, m, psi_l, out)  # and back: the slot epochs of both sequences stay in step
    e1f = float((np.abs(out[:, :rk.npw] - ref[:, rk.pw_l2g]).max(axis=1) / np.abs(ref).max(axis=1)).max())
    ok = ok and e1f < TOL
    print(f"rank {rank} R&G p2p unfused / fused again: hpsi err {e1u:.2e} {e1f:.2e}", flush=True)
    # sum_band density on the distributed descriptor (peer-to-peer exchange): this rank's z planes of the serial density
    rngw = np.random.default_rng(17)
    w = rngw.uniform(0.0, 2.0, m)
    rho_ref = R.sum_band_rho(d, inp["lda"], n, m, inp["psi"], w, np.zeros(d.nnr)).reshape(d.nr3, d.nr2 * d.nr1)
    z0 = int(np.sum(dd.nr3p[:rank]))
    nzl = int(dd.nr3p[rank])
    for rep in range(2):
        rho = np.zeros(d.nr1 * d.nr2 * nzl)
        hp.sum_band_rho(lda, rk.npw, m, np.nan_to_num(psi_l), w, rho)
        er = float(np.abs(rho.reshape(nzl, -1) - rho_ref[z0:z0 + nzl]).max() / np.abs(rho_ref).max())
        ok = ok and er < 1e-12
    print(f"rank {rank} sum_band density on the slab descriptor: err {er:.2e}", flush=True)
    out = np.full((m, lda), np.nan + 1j * np.nan)
    hp.h_psi(lda, rk.npw, m, psi_l, out)  # the V_loc exchange still lines up after the one-way density exchange
    e1d = float((np.abs(out[:, :rk.npw] - ref[:, rk.pw_l2g]).max(axis=1) / np.abs(ref).max(axis=1)).max())
    ok = ok and e1d < TOL
    dist.barrier()
    hp.close()
    # ---- band groups: replicated data, column blocks, NCCL all-gather inside the library ----
    hb = qe_b200.HPsi(device=local)
    hb.nccl_init(qe_b200.COMM_BGRP)
    hb.setup_from(d, inp)
    psi_t = torch.from_numpy(np.nan_to_num(inp["psi"])).cuda()
    out_t = torch.zeros_like(psi_t)
    hb.h_psi_bgrp(inp["lda"], n, m, psi_t, out_t, world, rank)
    torch.cuda.synchronize()
    e3 = colerr(out_t.cpu().numpy()[:, :n], ref[:, :n])
    ok = ok and e3 < TOL
    print(f"rank {rank} band groups: hpsi err {e3:.2e}", flush=True)
    hb.close()
    t = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0 and int(t.item()) == 1:
        print("DIST_OK", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if int(t.item()) == 1 else 1)


if __name__ == "__main__":
    main()
