#!/usr/bin/env python3
"""bench.py - H|psi> band-applications/s on B200 (BASELINE.json metric).

Workload (N=1): AUSURF112 (BASELINE configs[3]): smooth grid 125x64x200, npw = 95394,
nbnd = m = 800 bands (first Davidson call), nkb = 2016 ultrasoft projectors, one k-point.
A step = one h_psi(lda,n,m,psi,hpsi) call on synthetic random psi / V_loc / vkb / deeq.
N>1: k-point pools (one k-point per GPU, no collective inside the path) ->
"scaling": "weak"; `--mode bgrp` instead splits the m bands of ONE k-point over band groups
with an NCCL all-gather of hpsi (PW/src/h_psi.f90:53-67) -> "strong".

`--impl reference` times the reference's CPU algorithm for the same path on the host cores:
V_loc through the reference's own C FFT engine (oracle/_ref, compiled from
/root/reference/FFTXlib/src/fftw_dp.c, bands over OpenMP threads) and the projector GEMMs
through OpenBLAS (numpy), on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="ausurf112")
    ap.add_argument("--bands", type=int, default=0, help="override m (default nbnd of the workload)")
    ap.add_argument("--mode", default="pools", choices=["pools", "bgrp", "rg"])
    ap.add_argument("--cpu-sample-bands", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the output check of the timed step")
    ap.add_argument("--parity-cols", type=int, default=8)
    ap.add_argument("--no-library-bar", action="store_true", help="skip the cuFFT + cuBLAS bar")
    ap.add_argument("--no-davidson", action="store_true", help="skip the device-resident Davidson solve (qeb_cegterg)")
    ap.add_argument("--kpoints", type=int, default=16,
                    help="launch-bound workloads (si2, al4): k-points per step, submitted together with qeb_h_psi_multi")
    ap.add_argument("--no-strong", action="store_true", help="skip the strong-scaling modes measured beside the k-pools line")
    ap.add_argument("--davidson-ethr", type=float, default=1e-4)
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"],
                    help="--mode rg: stick<->plane exchange as peer-to-peer copy-engine writes (default) or NCCL send/recv")
    return ap.parse_args()


# ---------------------------------------------------------------------------
# clocks sampling (nvidia-smi during the timed region)
# ---------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.samples = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(self.index)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            f = [x.strip() for x in s.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------
# workload
# ---------------------------------------------------------------------------
def workload(name: str, seed: int = 0, impl: str = "ours"):
    """descriptor + sizes (host setup, not timed). Our arm uses the library's own descriptor builder (qeb_desc_*); the
    reference arm uses the oracle's (oracle/descriptor.py), so that process never maps libqeb200.so."""
    if impl == "reference":
        from oracle import descriptor as OD

        c = dict(OD.CONFIGS[name])
        c["nh"] = c.get("nh_list") or [c["nh"]] * c["nat"]
        d = OD.build_descriptor(c["at"], c["alat"], c["ecutwfc"], xk=c["xk"], gamma_only=c["gamma"], ecutrho=c["ecutrho"])
        return d, c, np.asarray(c["nh"], dtype=np.int32)
    import qe_b200

    c = dict(qe_b200.descriptor.WORKLOADS[name])
    xk = np.array(c["xk"], dtype=float)
    if not c["gamma"] and seed:
        xk = xk + 0.01 * seed  # a different k-point per pool
    d = qe_b200.descriptor.Descriptor(c["at"], c["alat"], c["ecutwfc"], c["ecutrho"], xk, c["gamma"])
    return d, c, np.asarray(c["nh"], dtype=np.int32)


def config_dict(workload_name, d, npw, nkb, m, n_gpus, mode="pools", exchange="p2p"):
    """the `config` object: identical keys and values in both arms for the same command line"""
    par = {"pools": "k-pools", "bgrp": "band-groups", "rg": "R&G slab FFT (" + exchange + " exchange)"}[mode]
    return {"workload": workload_name, "grid": [int(d.nr1), int(d.nr2), int(d.nr3)], "npw": int(npw), "nkb": int(nkb),
            "bands_per_step": int(m), "gamma_only": bool(d.gamma_only), "parallelism": f"{par} x{n_gpus}",
            "l2": "inputs larger than L2 (psi+hpsi+vkb = %.1f GB)" % ((2 * m * npw * 16 + nkb * npw * 16) / 1e9)}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return float(j.get("hbm_gbs", 6650.0)), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def cpu_reference_run(d, c, nh, m_sample: int, nthreads: int, reps: int = 1):
    """The reference's CPU algorithm on a bounded sample of m_sample bands: V_loc via the reference's
    C FFT engine (oracle/_ref), projectors via OpenBLAS zgemm/dgemm (numpy). Returns seconds per pass."""
    from oracle import hpsi_ref as R

    n = d.npw
    lda = n
    rng = np.random.default_rng(99)
    psi = (rng.standard_normal((m_sample, lda)) + 1j * rng.standard_normal((m_sample, lda))) / np.sqrt(2 * n)
    vrs = rng.uniform(-1, 0, d.nnr)
    nkb = int(nh.sum())
    vkb = np.empty((nkb, lda), dtype=np.complex128)
    for i0 in range(0, nkb, 256):  # chunked to bound temporary memory
        k = min(256, nkb - i0)
        vkb[i0:i0 + k] = (rng.standard_normal((k, lda)) + 1j * rng.standard_normal((k, lda))) / np.sqrt(n)
    nhm = int(nh.max())
    deeq = rng.uniform(-1, 1, (len(nh), nhm, nhm))
    ofs = np.concatenate([[0], np.cumsum(nh)[:-1]]).astype(np.int32)
    use_ref = R.ref_lib() is not None
    best = None
    # torchrun exports OMP_NUM_THREADS=1 to its workers: give OpenBLAS the host cores back for the reference's GEMMs
    # (the FFT engine gets its thread count explicitly)
    try:
        from threadpoolctl import threadpool_limits
        limiter = threadpool_limits(limits=nthreads)
    except Exception:  # noqa: BLE001
        limiter = None
    for _ in range(reps):
        t0 = time.perf_counter()
        hpsi = np.zeros((m_sample, lda), dtype=np.complex128)
        hpsi[:, :n] = d.g2kin[None, :n] * psi[:, :n]
        if use_ref:
            R.ref_vloc_psi(d, lda, n, m_sample, psi, vrs, hpsi, nthreads)
        else:
            (R.vloc_psi_gamma if d.gamma_only else R.vloc_psi_k)(d, lda, n, m_sample, psi, vrs, hpsi)
        if d.gamma_only:
            becp = R.calbec_gamma(n, vkb, psi, m_sample, d.gstart)
            ps = R.apply_deeq(becp, deeq, nh, ofs)
            R.add_vuspsi_gamma(n, m_sample, vkb, ps, hpsi)
        else:
            becp = R.calbec_k(n, vkb, psi, m_sample)
            ps = R.apply_deeq(becp, deeq, nh, ofs)
            R.add_vuspsi_k(n, m_sample, vkb, ps, hpsi)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    if limiter is not None:
        limiter.restore_original_limits()
    return best, ("reference" if use_ref else "port")


def parity_check(torch, d, lda, n, m, psi, hpsi, vkb, vrs, deeq, nh, ofs, k=8):
    """k evenly spaced columns of the step's own output against the checker: the code the reference arm times
    (oracle: reference C FFT engine for V_loc when built, numpy/OpenBLAS zgemm restating becmod.f90:431-486 and
    add_vuspsi.f90:140-170). Per-column max|d| / max|ref column|."""
    from oracle import hpsi_ref as R  # checker only

    cols = np.unique(np.linspace(0, m - 1, min(k, m)).astype(np.int64))
    if d.gamma_only:  # whole band pairs, as the Gamma path transforms them
        cols = np.unique(np.concatenate([cols - cols % 2, np.minimum(cols - cols % 2 + 1, m - 1)]))
    ct = torch.from_numpy(cols).to(psi.device)
    ps = np.ascontiguousarray(torch.view_as_complex(psi[ct].contiguous()).cpu().numpy())
    got = torch.view_as_complex(hpsi[ct].contiguous()).cpu().numpy()
    kk = len(cols)
    ref = np.zeros((kk, lda), dtype=np.complex128)
    ref[:, :n] = d.g2kin[None, :n] * ps[:, :n]
    use_ref = R.ref_lib() is not None
    if use_ref:
        R.ref_vloc_psi(d, lda, n, kk, ps, vrs, ref, os.cpu_count() or 1)
    else:
        (R.vloc_psi_gamma if d.gamma_only else R.vloc_psi_k)(d, lda, n, kk, ps, vrs, ref)
    nkb = int(np.sum(nh))
    if nkb:
        vk = torch.view_as_complex(vkb).cpu().numpy()
        if d.gamma_only:
            becp = R.calbec_gamma(n, vk, ps, kk, d.gstart)
            R.add_vuspsi_gamma(n, kk, vk, R.apply_deeq(becp, deeq, nh, ofs), ref)
        else:
            becp = R.calbec_k(n, vk, ps, kk)
            R.add_vuspsi_k(n, kk, vk, R.apply_deeq(becp, deeq, nh, ofs), ref)
        del vk
    if d.gamma_only and d.gstart == 2:
        ref[:, 0] = ref[:, 0].real
    num = np.abs(got[:, :n] - ref[:, :n]).max(axis=1)
    den = np.abs(ref[:, :n]).max(axis=1)
    return {"cols": int(kk), "max_rel": float((num / den).max()), "tolerance": 1e-11,
            "checker": ("reference C FFT engine (oracle/_ref)" if use_ref else "numpy restatement") + " + numpy zgemm",
            "pad_rows_zero": bool(np.all(got[:, n:] == 0))}


# ---------------------------------------------------------------------------
# strong-scaling modes, measured in the same run as the k-pools line
# ---------------------------------------------------------------------------
def strong_mode(torch, dist, qe_b200, wname, mode, world, rank, local, stream, steps=3, group=None, gsize=None, grank=None,
                exchange="p2p", seed=0):
    """one h_psi call of `wname` per step with the work of ONE k-point split over `gsize` ranks (the ranks of `group`):
    mode "bgrp": band groups + all-gather of hpsi (h_psi.f90:53-67); mode "rg": slab-distributed FFT + becp all-reduce
    (fft_parallel_2d.f90, becmod.f90:395-399). Returns (bands per step of the group, ms per step as max over ranks)."""
    gsize = gsize or world
    grank = rank if grank is None else grank
    dev = torch.device("cuda", local)
    d, c, nh = workload(wname, seed=seed)
    m_total = c["nbnd"]
    nkb = int(nh.sum())
    rg = mode == "rg"
    if rg:
        d.distribute(gsize)
        rk = d.ranks[grank]
        n = rk.npw
    else:
        n = d.npw
    lda = n
    g = torch.Generator(device=dev).manual_seed(99 + grank)
    psi = torch.randn((m_total, lda, 2), generator=g, device=dev, dtype=torch.float64) / np.sqrt(2 * d.npw)
    if d.gamma_only:
        psi[:, 0, 1] = 0.0
    hpsi = torch.zeros_like(psi)
    vkb = torch.randn((nkb, lda, 2), generator=g, device=dev, dtype=torch.float64) / np.sqrt(d.npw)
    if d.gamma_only:
        vkb[:, 0, 1] = 0.0
    rng = np.random.default_rng(77 + grank)
    nhm = int(nh.max())
    deeq = np.random.default_rng(4234).uniform(-1, 1, (len(nh), nhm, nhm))
    deeq = 0.5 * (deeq + deeq.transpose(0, 2, 1))
    ofs = np.concatenate([[0], np.cumsum(nh)[:-1]]).astype(np.int32)
    h = qe_b200.HPsi(device=local, stream=stream)
    if rg:
        h.set_grid_dist(d.nr1, d.nr2, d.nr3, gsize, grank, d.nr3p, d.nsw, d.iss, d.ismap, rk.nl,
                        rk.nlm if d.gamma_only else None, d.gamma_only)
        if gsize > 1:
            h.nccl_init(qe_b200.COMM_RG)
        h.set_kpoint(n, lda, None if d.gamma_only else rk.igk, rk.g2kin, vkb, nkb, rk.gstart)
        h.set_vloc(rng.uniform(-1.0, 0.0, d.nr1 * d.nr2 * int(d.nr3p[grank])))
        if gsize > 1 and exchange == "p2p":
            h.p2p_setup()
    else:
        h.set_grid(d.nr1, d.nr2, d.nr3, d.nl, d.nlm if d.gamma_only else None, d.gamma_only)
        h.set_kpoint(n, lda, None if d.gamma_only else d.igk, d.g2kin, vkb, nkb, d.gstart)
        h.set_vloc(rng.uniform(-1.0, 0.0, d.nnr))
    h.set_nonlocal(nh, ofs, deeq)
    h.set_async(True)
    m0, m1 = (0, m_total) if rg else qe_b200.divide_all(m_total, gsize, grank)
    bounds = [qe_b200.divide_all(m_total, gsize, r) for r in range(gsize)]

    def step():
        if rg or gsize == 1:
            h.h_psi(lda, n, m_total, psi, hpsi)
            return
        h.h_psi_bgrp(lda, n, m_total, psi, hpsi, gsize, grank)
        blocks = [hpsi[s_:e_] for s_, e_ in bounds]
        if len({b.shape[0] for b in blocks}) == 1:
            dist.all_gather(blocks, hpsi[m0:m1], group=group)
        else:
            for r, b in enumerate(blocks):
                dist.broadcast(b, src=dist.get_global_rank(group, r) if group is not None else r, group=group)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(3):
        step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1) / steps], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    h.close()
    del psi, hpsi, vkb
    torch.cuda.empty_cache()
    return m_total, float(t.item())


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    d, c, nh = workload(args.workload, impl="reference")
    cores = os.cpu_count() or 1
    m_s = args.cpu_sample_bands or args.bands or int(c.get("nbnd", 128))
    times = []
    kind = "port"
    for i in range(args.warmup + args.steps):
        dt, kind = cpu_reference_run(d, c, nh, m_s, cores)
        if i >= args.warmup:
            times.append(dt)
    t = float(np.mean(times))
    val = m_s / t
    line = {
        "impl": "reference", "metric": "h_psi band-applications/s", "value": val, "unit": "band-applications/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_dict(args.workload, d, d.npw, int(nh.sum()), m_s, args.gpus, args.mode, args.exchange),
        "cpu_baseline": {"value": val, "unit": "band-applications/s", "cores": cores, "kind": kind,
                         "sample": f"{m_s} bands of the {args.workload} h_psi: V_loc via the reference C FFT engine "
                                   f"(OpenMP over bands), projectors via OpenBLAS zgemm"},
        "e2e": {"value": val, "unit": "band-applications/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    # this arm must not map the product library (only oracle/_ref + numpy/OpenBLAS)
    try:
        line["maps_libqeb200"] = "libqeb200" in open("/proc/self/maps").read()
    except OSError:
        line["maps_libqeb200"] = None
    emit(line)


def run_ours(args):
    import torch
    import torch.distributed as dist

    import qe_b200

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: a CUDA device is required (no CPU fallback)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)

    pools = args.mode == "pools"
    rg = args.mode == "rg"
    d, c, nh = workload(args.workload, seed=rank if pools else 0)
    m_total = args.bands or c["nbnd"]
    nkb = int(nh.sum())
    if rg:
        # R&G: sticks / planes / plane waves of ONE k-point distributed over the ranks
        d.distribute(world)
        rk = d.ranks[rank]
        n = rk.npw
    else:
        n = d.npw
    lda = n
    if pools or rg:
        m0, m1 = 0, m_total
    else:
        m0, m1 = qe_b200.divide_all(m_total, world, rank)
    m_loc = m1 - m0

    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    psi = torch.randn((m_total, lda, 2), generator=g, device=dev, dtype=torch.float64) / np.sqrt(2 * d.npw)
    if d.gamma_only:
        psi[:, 0, 1] = 0.0
    hpsi = torch.zeros_like(psi)
    vkb = torch.randn((nkb, lda, 2), generator=g, device=dev, dtype=torch.float64) / np.sqrt(d.npw)
    if d.gamma_only:
        vkb[:, 0, 1] = 0.0
    rng = np.random.default_rng(2234 + rank)
    nhm = int(nh.max())
    deeq = np.random.default_rng(4234).uniform(-1, 1, (len(nh), nhm, nhm))
    deeq = 0.5 * (deeq + deeq.transpose(0, 2, 1))
    ofs = np.concatenate([[0], np.cumsum(nh)[:-1]]).astype(np.int32)

    # a dedicated (non-default) stream: the library launches on it and the CUDA events below time it
    tstream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(tstream)
    stream = tstream.cuda_stream
    assert stream != 0
    h = qe_b200.HPsi(device=local, stream=stream, batch=args.batch or None)
    if rg:
        h.set_grid_dist(d.nr1, d.nr2, d.nr3, world, rank, d.nr3p, d.nsw, d.iss, d.ismap, rk.nl,
                        rk.nlm if d.gamma_only else None, d.gamma_only)
        if world > 1:
            h.nccl_init(qe_b200.COMM_RG)
        h.set_kpoint(n, lda, None if d.gamma_only else rk.igk, rk.g2kin, vkb, nkb, rk.gstart)
        vrs = rng.uniform(-1.0, 0.0, d.nr1 * d.nr2 * int(d.nr3p[rank]))
        h.set_vloc(vrs)
        if world > 1 and args.exchange == "p2p":
            h.p2p_setup()
    else:
        h.set_grid(d.nr1, d.nr2, d.nr3, d.nl, d.nlm if d.gamma_only else None, d.gamma_only)
        h.set_kpoint(n, lda, None if d.gamma_only else d.igk, d.g2kin, vkb, nkb, d.gstart)
        vrs = rng.uniform(-1.0, 0.0, d.nnr)
        h.set_vloc(vrs)
    h.set_nonlocal(nh, ofs, deeq)
    h.set_async(True)
    st = h.stats()
    # launch-bound workloads (BASELINE configs 1-2: a few bands on a 25^3 grid): the k-points of the pool are submitted
    # together (qeb_h_psi_multi: one context and stream per k-point, CUDA-graph replay), as the k loop of c_bands would
    multi = []
    if pools and not rg and d.npw * m_total <= (1 << 17) and args.kpoints > 1:
        for kk in range(1, args.kpoints):
            dk, _, _ = workload(args.workload, seed=kk + args.kpoints * rank)
            nk_ = dk.npw
            gk = torch.Generator(device=dev).manual_seed(4321 + kk)
            pk = torch.randn((m_total, nk_, 2), generator=gk, device=dev, dtype=torch.float64) / np.sqrt(2 * nk_)
            vk = torch.randn((nkb, nk_, 2), generator=gk, device=dev, dtype=torch.float64) / np.sqrt(nk_)
            hk = qe_b200.HPsi(device=local)
            hk.set_grid(dk.nr1, dk.nr2, dk.nr3, dk.nl, None, False)
            hk.set_kpoint(nk_, nk_, dk.igk, dk.g2kin, vk, nkb, dk.gstart)
            hk.set_vloc(vrs)
            hk.set_nonlocal(nh, ofs, deeq)
            multi.append((hk, nk_, pk, torch.zeros_like(pk), vk))
    nk_step = 1 + len(multi)
    if multi:
        m_ctx = [h] + [t_[0] for t_ in multi]
        m_lda = [lda] + [t_[1] for t_ in multi]
        m_psi = [psi] + [t_[2] for t_ in multi]
        m_out = [hpsi] + [t_[3] for t_ in multi]

    def step():
        if multi:
            qe_b200.h_psi_multi(m_ctx, m_lda, m_lda, [m_total] * nk_step, m_psi, m_out)
        elif pools or rg or world == 1:
            h.h_psi(lda, n, m_total, psi, hpsi)
        else:
            h.h_psi_bgrp(lda, n, m_total, psi, hpsi, world, rank)
            # all-gather of the column blocks over inter_bgrp_comm (h_psi.f90:58-63)
            blocks = [hpsi[s:e] for s, e in (qe_b200.divide_all(m_total, world, r) for r in range(world))]
            dist.all_gather(blocks, hpsi[m0:m1]) if len({b.shape[0] for b in blocks}) == 1 else [
                dist.broadcast(b, src=r) for r, b in enumerate(blocks)]

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # (multi-k graphs: the first call sizes buffers, the second sees the final key, the third captures)
    for _ in range(max(args.warmup, 3) + (2 if multi else 0)):
        step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    clocks = sampler.stop() if rank == 0 else None
    ms_step = ms_total / args.steps
    units = m_total * (world if pools else 1) * nk_step
    value = units / (ms_step * 1e-3)

    # ---- per-section timing on the launching stream (V_loc alone, projectors alone) ----
    def time_section(fn, reps=3):
        fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    mm = m_total if (pools or rg) else m_loc
    ms_vloc = time_section(lambda: (h.vloc_psi_gamma if d.gamma_only else h.vloc_psi_k)(lda, n, mm, psi, None, hpsi))
    ms_nl = time_section(lambda: h.nonlocal_only(lda, n, mm, psi, hpsi)) if nkb else 0.0
    peak, peak_src = peaks()
    boxes = (mm + 1) // 2 if d.gamma_only else mm
    nbatches = -(-boxes // st["batch"])
    vloc_bytes = st["vloc_bytes_per_band"] * mm
    achieved = vloc_bytes / (ms_vloc * 1e-3) / 1e9
    # V_loc pipeline (5 pass kernels per batch of boxes) against the HBM roofline
    roofline_vloc = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": None, "kernel": "V_loc pass pipeline (5 kernels/batch: zinv,yinv,xfus,yfwd,zfwd)",
                     "algorithmic_bytes_per_band": st["vloc_bytes_per_band"], "ms_vloc_section": ms_vloc,
                     "peak_source": peak_src}
    gemm_tf_alg = st["gemm_flops_per_band"] * mm / (ms_nl * 1e-3) / 1e12 if ms_nl > 0 else None
    # complex (k-point) GEMMs run in the 3-multiplication form: 3 real MMAs per complex multiply-add, so the DMMA pipe
    # executes 12*nkb*npw flops per band for the 16*nkb*npw algorithmic ones; the roofline fraction uses the executed count
    three_m = (not d.gamma_only) and os.environ.get("QEB_GEMM_4M", "0") in ("", "0")
    gemm_tf = gemm_tf_alg * (0.75 if three_m else 1.0) if gemm_tf_alg else None
    # FP64 tensor (DMMA) ceiling: 148 SMs x 64 FMA/clk x SM clock; cuBLAS DGEMM timed beside it as the empirical ceiling
    sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0
    fp64_peak = torch.cuda.get_device_properties(dev).multi_processor_count * 64 * 2 * sm_mhz * 1e6 / 1e12
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import library_bar as LB  # bench infrastructure (cuFFT / cuBLAS through torch), not part of libqeb200

    fp64 = LB.fp64_ceiling(torch, dev) if rank == 0 else {"dgemm_8192_tflops": float("nan"), "zgemm_8192_tflops": float("nan")}
    fp64["dmma_peak_tflops"] = fp64_peak
    fp64["how"] = ("cuBLAS DGEMM / ZGEMM 8192^3 (torch.matmul, 3 reps, CUDA events) in this run; dmma_peak = 148 SMs x 64 "
                   "FP64 FMA/clk x the SM clock sampled during the timed region")
    dgemm_tf = fp64["dgemm_8192_tflops"]
    roofline_gemm = None
    if ms_nl > 0:
        roofline_gemm = {"bound": "tensor", "achieved": gemm_tf, "peak": fp64_peak, "unit": "TFLOP/s",
                         "frac": gemm_tf / fp64_peak, "traffic": None,
                         "kernel": "projector GEMMs k_calbec + k_addvus (FP64 DMMA.8x8x4)",
                         "algorithmic_flops_per_band": st["gemm_flops_per_band"], "algorithmic_tflops": gemm_tf_alg,
                         "complex_product": "3M (executed flops = 0.75 x algorithmic)" if three_m else "4M / real",
                         "ms_projector_section": ms_nl,
                         "peak_source": "148 SMs x 64 FP64 FMA/clk x %.0f MHz (FP64 tensor = FP64 vector rate on B200); "
                                        "cuBLAS DGEMM 8192^3 in this run: %.1f TFLOP/s, ZGEMM %.1f" % (sm_mhz, dgemm_tf, fp64["zgemm_8192_tflops"])}
    kpb = h.vloc_kernels_per_batch()  # 3 with the fused plane kernel (zinv, plane, zfwd), else 5
    # DRAM traffic (dram__bytes_read + write per launch) of the same kernels from the committed `ncu --set full` summary
    # profiles/traffic.json, written by tools/ncu_summary.py from the capture of the commit it names
    tj = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tj):
        tr = json.load(open(tj)).get(args.workload, {})
        if roofline_gemm and three_m and tr.get("gemm_bytes_per_band"):
            roofline_gemm["traffic"] = tr["gemm_bytes_per_band"] * mm
            roofline_gemm["traffic_source"] = tr.get("gemm_source")
        key = "vloc_bytes_per_box_plane" if kpb == 3 else "vloc_bytes_per_box_5k"
        if tr.get(key):
            roofline_vloc["traffic"] = tr[key] * boxes
            roofline_vloc["traffic_source"] = tr.get("vloc_source")
    # `roofline` = the kernel family with the largest share of the step
    roofline = roofline_gemm if (roofline_gemm and ms_nl >= ms_vloc) else roofline_vloc
    launches_step = nbatches * kpb + (4 if nkb else 0) + 1
    roofline_vloc["kernel"] = ("V_loc pass pipeline (3 kernels/batch: zinv, fused plane y|x*V*x|y, zfwd)" if kpb == 3 else
                               "V_loc pass pipeline (5 kernels/batch: zinv,yinv,xfus,yfwd,zfwd)")

    # ---- end to end through the public API with HOST (pinned) buffers ----
    e2e = None
    if not args.no_e2e:
        h.set_async(False)
        psi_h = torch.empty((mm, lda, 2), dtype=torch.float64).pin_memory()
        psi_h.copy_(psi[:mm].cpu())
        hpsi_h = torch.empty((mm, lda, 2), dtype=torch.float64).pin_memory()
        pn, hn = psi_h.numpy(), hpsi_h.numpy()
        h.h_psi(lda, n, mm, pn, hn)
        barrier()
        t0 = time.perf_counter()
        reps = max(5, args.steps)
        for _ in range(reps):
            h.h_psi(lda, n, mm, pn, hn)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / reps
        tt = torch.tensor([dt], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
        e2e = {"value": (mm * world if pools else m_total) / dt, "unit": "band-applications/s",
               "h2d_bytes_per_step": int(mm * lda * 16), "d2h_bytes_per_step": int(mm * lda * 16),
               "ms_per_step": dt * 1e3}
        h.set_async(True)

    # ---- the path inside its caller: one device-resident block-Davidson solve (qeb_cegterg / qeb_regterg, cegterg.f90) ----
    # psi/hpsi/spsi stay in HBM for the whole solve; the host sees evc once in and once out (pinned), the eigenvalues and
    # the counters. Rate = H applications of the solve / wall time of the call (S, the reduced-matrix GEMMs, the basis
    # updates and the nbase x nbase cuSOLVER eigenproblems included), max over ranks; k-pools mode: one solve per rank.
    davidson = None
    if not args.no_davidson and not rg and (pools or world == 1) and nkb:
        nvec, nvecx = mm, 2 * mm
        qrng = np.random.default_rng(5234)
        qq = np.zeros((len(nh), nhm, nhm))
        for a_, k_ in enumerate(nh):
            A_ = qrng.uniform(-1, 1, (k_, k_))
            qq[a_, :k_, :k_] = 0.05 * (A_ @ A_.T)  # positive semi-definite: S = 1 + sum |beta> qq <beta| > 0
        h.set_qq(qq)
        g2 = np.asarray((rk.g2kin if rg else d.g2kin)[:n], dtype=np.float64)
        h.set_g_psi_diag(g2 + float(np.mean(vrs)), np.ones(n))
        evc_h = torch.empty((nvec, lda, 2), dtype=torch.float64).pin_memory()
        evc_h.copy_(psi[:nvec].cpu())
        btype = np.ones(nvec, dtype=np.int32)
        h.set_async(False)
        try:
            barrier()
            t0 = time.perf_counter()
            ev, notcnv, dav_iter, nhpsi = h.cegterg(n, lda, nvec, nvecx, evc_h.numpy(), args.davidson_ethr, True, btype)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            agg = torch.tensor([dt, float(nhpsi)], device=dev, dtype=torch.float64)
            if world > 1:
                mx = agg.clone()
                dist.all_reduce(mx, op=dist.ReduceOp.MAX)
                dist.all_reduce(agg, op=dist.ReduceOp.SUM)
                dt_max, nh_sum = float(mx[0].item()), float(agg[1].item())
            else:
                dt_max, nh_sum = dt, float(nhpsi)
            davidson = {"entry": "qeb_regterg" if d.gamma_only else "qeb_cegterg", "nvec": nvec, "nvecx": nvecx,
                        "ethr": args.davidson_ethr, "uspp": True, "dav_iter": dav_iter, "notcnv": notcnv, "nhpsi": nhpsi,
                        "ms": dt_max * 1e3, "value": nh_sum / dt_max, "unit": "band-applications/s",
                        "h2d_bytes_per_solve": int(nvec * lda * 16), "d2h_bytes_per_solve": int(nvec * lda * 16),
                        "host_bytes_per_h_application": 2.0 * nvec * lda * 16 / max(nhpsi, 1),
                        "e_min": float(ev.min()), "e_max": float(ev.max())}
        except Exception as exc:  # noqa: BLE001  (e.g. cuSOLVER not found): reported, never fatal
            davidson = {"unavailable": repr(exc)[:300]}
        h.set_async(True)

    # ---- output check of the timed step: a sample of columns against the checker (the reference arm's code) ----
    parity = None
    lib_bar = None
    if rank == 0 and not rg and not args.no_parity:
        step()  # the timed step once more (hpsi was overwritten by the section timings)
        torch.cuda.synchronize()
        parity = parity_check(torch, d, lda, n, mm, psi, hpsi, vkb, vrs, deeq, nh, ofs, k=args.parity_cols)
    # ---- the same-box GPU library bar: cuFFT full-box Z2Z + cuBLAS ZGEMM composed as the reference's GPU path ----
    if rank == 0 and world == 1 and not rg and not args.no_library_bar:
        try:
            lib_bar = LB.measure(torch, d, psi[:mm], vkb, vrs, deeq, nh, ofs, reps=2)
            Hl = lib_bar.pop("hpsi")
            ours = torch.view_as_complex(hpsi[:mm])
            if d.gamma_only and d.gstart == 2:
                Hl[:, 0] = Hl[:, 0].real
            num = (Hl[:, :n] - ours[:, :n]).abs().amax(dim=1)
            den = Hl[:, :n].abs().amax(dim=1)
            lib_bar["max_rel_vs_ours"] = float((num / den).max().item())  # per column, all bands of the step
            lib_bar["ours_over_library"] = value / lib_bar["value"]
            lib_bar["vloc_ours_over_library"] = lib_bar["vloc_us_per_band"] / (1e3 * ms_vloc / mm)
            lib_bar["projectors_ours_over_library"] = (lib_bar["projectors_us_per_band"] / (1e3 * ms_nl / mm)) if ms_nl else None
            del Hl
        except Exception as exc:  # noqa: BLE001  (e.g. out of memory on a shared box): reported, never fatal
            lib_bar = {"unavailable": repr(exc)[:200]}
        torch.cuda.empty_cache()

    # ---- strong-scaling modes of ONE k-point over the same GPUs (values only; efficiencies follow from the N=1 run) ----
    strong = None
    if pools and not args.no_strong and args.workload == "ausurf112":
        strong = {}
        try:
            mb, msb = strong_mode(torch, dist, qe_b200, "ausurf112", "bgrp", world, rank, local, stream)
            strong["bgrp"] = {"workload": "ausurf112", "layout": f"{world} band groups", "bands_per_step": mb, "ms_per_step": msb,
                              "value": mb / (msb * 1e-3), "unit": "band-applications/s"}
            if world >= 4 and world % 2 == 0:
                half = world // 2
                groups = [dist.new_group(list(range(p_ * half, (p_ + 1) * half))) for p_ in range(2)]
                pool = rank // half
                mb, msb = strong_mode(torch, dist, qe_b200, "ausurf112", "bgrp", world, rank, local, stream, group=groups[pool],
                                      gsize=half, grank=rank % half, seed=pool)
                strong["pools2xbgrp"] = {"workload": "ausurf112", "layout": f"2 k-pools x {half} band groups (the 2 k-points of AUSURF112)",
                                         "bands_per_step": 2 * mb, "ms_per_step": msb, "value": 2 * mb / (msb * 1e-3),
                                         "unit": "band-applications/s"}
            mb, msb = strong_mode(torch, dist, qe_b200, "si512", "rg", world, rank, local, stream, exchange=args.exchange)
            strong["rg_si512"] = {"workload": "si512", "layout": f"R&G slab FFT over {world} ranks ({args.exchange} exchange)",
                                  "bands_per_step": mb, "ms_per_step": msb, "value": mb / (msb * 1e-3), "unit": "band-applications/s"}
        except Exception as exc:  # noqa: BLE001
            strong["error"] = repr(exc)[:300]

    if rank == 0:
        line = {
            "metric": "h_psi band-applications/s", "value": value, "unit": "band-applications/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak" if pools else "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_dict(args.workload, d, d.npw, nkb, m_total, world, args.mode, args.exchange),
            "tuning": {"vloc_batch_boxes": st["batch"], "vloc_kernels_per_batch": kpb, "k_points_per_step": nk_step,
                       "cuda_graph": bool(multi)},
            "fp64_ceiling": fp64,
            "roofline": roofline,
            "roofline_vloc": roofline_vloc,
            "roofline_projectors": roofline_gemm,
            "sections_ms": {"vloc": ms_vloc, "projectors": ms_nl},
            "parity_check": parity,
            "gpu_library_bar": lib_bar,
            "davidson": davidson,
            "strong": strong,
            "gpu_launches": launches_step * args.steps,
            "clocks": clocks,
        }
        if e2e:
            line["e2e"] = e2e
        if not args.no_cpu_baseline and world == 1:
            cores = os.cpu_count() or 1
            m_s = args.cpu_sample_bands or min(int(c.get("nbnd", 128)), 400)
            dt, kind = cpu_reference_run(d, c, nh, m_s, cores)
            line["cpu_baseline"] = {"value": m_s / dt, "unit": "band-applications/s", "cores": cores, "kind": kind,
                                    "sample": f"{m_s} bands of the same h_psi (V_loc: reference C FFT engine, OpenMP "
                                              f"over bands; projectors: OpenBLAS zgemm), {dt:.1f} s"}
        emit(line)
    h.close()
    for t_ in multi:
        t_[0].close()
    if world > 1:
        dist.destroy_process_group()


_REAL_STDOUT = None


def emit(line: dict):
    """the ONE JSON line, on the process's real stdout"""
    data = (json.dumps(line) + "\n").encode()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, data)


def main():
    global _REAL_STDOUT
    # libraries (NCCL prints its version banner on stdout) must not pollute the one-line contract: everything
    # written to fd 1 while the benchmark runs goes to stderr, the JSON line goes to the saved stdout
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
