"""ctypes mirror of include/qeb200.h with the reference's names and argument
meaning (PW/src/h_psi.f90:9, PW/src/vloc_psi_acc.f90:9,148,
FFTXlib/src/fft_interfaces.f90:19-80, Modules/becmod.f90:138,
PW/src/add_vuspsi.f90:10). Arrays follow the reference's memory layout: a
Fortran psi(lda, m) is a C-contiguous numpy/torch array of shape (m, lda).
Arguments may be numpy arrays (host) or torch CUDA tensors (device).
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

FFT_RHO = 0
FFT_WAVE = 1
COMM_BGRP = 0  # inter_bgrp_comm (band groups)
COMM_RG = 1    # intra_bgrp_comm = dffts%comm (R&G distribution)


class QebError(RuntimeError):
    pass


def lib_path() -> str:
    # QEB_LIB selects another build of the same library (kernel-tuning variants, q-e_b200/build.py --variant)
    return os.environ.get("QEB_LIB") or os.path.join(_HERE, "libqeb200.so")


def load_library():
    """Loads libqeb200.so (building it is __graft_entry__.build()'s job)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = lib_path()
    if not os.path.exists(path):
        raise QebError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` (no CPU fallback)")
    L = ctypes.CDLL(path)
    P, I = ctypes.c_void_p, ctypes.c_int
    L.qeb_create.argtypes = [ctypes.POINTER(P), I, P]
    L.qeb_destroy.argtypes = [P]
    L.qeb_last_error.argtypes = [P]
    L.qeb_last_error.restype = ctypes.c_char_p
    L.qeb_version.restype = ctypes.c_char_p
    L.qeb_set_batch.argtypes = [P, I]
    L.qeb_set_async.argtypes = [P, I]
    L.qeb_set_pipeline.argtypes = [P, I, I]
    L.qeb_pin_host.argtypes = [P, P, ctypes.c_size_t]
    L.qeb_unpin_host.argtypes = [P, P]
    L.qeb_device_count.argtypes = []
    L.qeb_set_grid.argtypes = [P, I, I, I, I, I, I, I, P, P, I]
    L.qeb_set_kpoint.argtypes = [P, I, I, P, P, P, I, I]
    L.qeb_set_vloc.argtypes = [P, P]
    L.qeb_set_nonlocal.argtypes = [P, I, I, P, P, P]
    L.qeb_h_psi.argtypes = [P, I, I, I, P, P]
    L.qeb_h_psi_multi.argtypes = [I, P, P, P, P, P, P]
    L.qeb_hs_psi.argtypes = [P, I, I, I, P, P, P]
    L.qeb_cegterg.argtypes = [P, I, I, I, I, I, P, ctypes.c_double, I, P, P, P, I, P, P]
    L.qeb_regterg.argtypes = [P, I, I, I, I, P, ctypes.c_double, I, P, P, P, I, P, P]
    L.qeb_vloc_psi_k.argtypes = [P, I, I, I, P, P, P]
    L.qeb_vloc_psi_gamma.argtypes = [P, I, I, I, P, P, P]
    L.qeb_set_vloc_nc.argtypes = [P, P]
    L.qeb_vloc_psi_nc.argtypes = [P, I, I, I, P, P, P]
    L.qeb_set_nonlocal_nc.argtypes = [P, P]
    L.qeb_h_psi_nc.argtypes = [P, I, I, I, P, P]
    L.qeb_set_qq_nc.argtypes = [P, P]
    L.qeb_s_psi_nc.argtypes = [P, I, I, I, P, P]
    L.qeb_fft.argtypes = [P, I, I, P, I]
    L.qeb_calbec.argtypes = [P, I, P, I, P, I, I, P]
    L.qeb_calbec_static.argtypes = [P, I, P, I, P, I, I, P]
    L.qeb_calbec_static.restype = I
    L.qeb_static_scratch_bytes.argtypes = []
    L.qeb_static_scratch_bytes.restype = ctypes.c_size_t
    L.qeb_add_vuspsi.argtypes = [P, I, I, I, P, P]
    L.qeb_get_becp.argtypes = [P, P, I]
    L.qeb_set_qq.argtypes = [P, I, I, P]
    L.qeb_s_psi.argtypes = [P, I, I, I, P, P]
    D = ctypes.c_double
    L.qeb_set_gvectors.argtypes = [P, I, P, P]
    L.qeb_g2_kin.argtypes = [P, P, D, D, D, D, P]
    L.qeb_set_beta_tables.argtypes = [P, I, I, I, I, P, P, P, P, I, D, P]
    L.qeb_set_structure.argtypes = [P, I, P, P, I, I, I, P, P, P]
    L.qeb_init_us_2.argtypes = [P, P, D, P]
    L.qeb_sum_band_rho.argtypes = [P, I, I, I, P, P, P]
    L.qeb_sum_band_rho_nc.argtypes = [P, I, I, I, P, P, P, I]
    L.qeb_sum_band_rho_nc.restype = I
    L.qeb_p2p_export.argtypes = [P, P]
    L.qeb_p2p_import.argtypes = [P, P, I]
    L.qeb_set_g_psi_diag.argtypes = [P, I, P, P]
    L.qeb_g_psi.argtypes = [P, I, I, I, I, P, P]
    L.qeb_col_dot.argtypes = [P, I, I, P, I, P, I, P]
    L.qeb_dav_gemm_cn.argtypes = [P, I, P, I, I, P, I, I, P, I]
    L.qeb_dav_gemm_nn.argtypes = [P, I, P, I, I, P, I, I, P, I, I]
    L.qeb_divide_all.argtypes = [I, I, I, ctypes.POINTER(I), ctypes.POINTER(I)]
    L.qeb_h_psi_bgrp.argtypes = [P, I, I, I, P, P, I, I]
    L.qeb_set_nccl.argtypes = [P, P, I, I]
    L.qeb_set_nccl_comm.argtypes = [P, I, P, I, I]
    L.qeb_nccl_unique_id.argtypes = [P]
    L.qeb_nccl_init.argtypes = [P, I, P, I, I]
    L.qeb_set_grid_dist.argtypes = [P, I, I, I, I, I, P, P, P, P, I, I, P, P, I]
    L.qeb_get_stats.argtypes = [P, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double),
                                ctypes.POINTER(I), ctypes.POINTER(I), ctypes.POINTER(I)]
    L.qeb_nonlocal_only.argtypes = [P, I, I, I, P, P]
    L.qeb_vloc_kernels_per_batch.argtypes = [P]
    for name in ("qeb_create", "qeb_destroy", "qeb_device_count", "qeb_pin_host", "qeb_unpin_host", "qeb_set_pipeline", "qeb_set_batch", "qeb_set_async", "qeb_set_grid", "qeb_set_kpoint",
                 "qeb_set_vloc", "qeb_set_nonlocal", "qeb_h_psi", "qeb_vloc_psi_k", "qeb_vloc_psi_gamma", "qeb_set_vloc_nc", "qeb_vloc_psi_nc", "qeb_set_nonlocal_nc", "qeb_h_psi_nc", "qeb_set_qq_nc", "qeb_s_psi_nc", "qeb_fft",
                 "qeb_calbec", "qeb_add_vuspsi", "qeb_get_becp", "qeb_divide_all", "qeb_h_psi_bgrp", "qeb_set_nccl",
                 "qeb_get_stats", "qeb_nonlocal_only", "qeb_set_nccl_comm", "qeb_nccl_unique_id", "qeb_nccl_init",
                 "qeb_set_grid_dist", "qeb_set_qq", "qeb_s_psi", "qeb_set_gvectors", "qeb_g2_kin", "qeb_set_beta_tables",
                 "qeb_set_structure", "qeb_init_us_2", "qeb_set_g_psi_diag", "qeb_g_psi", "qeb_col_dot",
                 "qeb_dav_gemm_cn", "qeb_dav_gemm_nn", "qeb_sum_band_rho", "qeb_p2p_export", "qeb_p2p_import"):
        getattr(L, name).restype = I
    _LIB = L
    return L


def divide_all(m: int, nproc: int, me: int):
    """UtilXlib/divide.f90:55-106 -> 0-based [start, end) of group `me`."""
    L = load_library()
    s, e = ctypes.c_int(), ctypes.c_int()
    if L.qeb_divide_all(m, nproc, me, ctypes.byref(s), ctypes.byref(e)):
        raise QebError("divide_all: bad arguments")
    return s.value, e.value


def allgather_columns(hpsi, m: int, group=None):
    """The band-group all-gather of PW/src/h_psi.f90:58-63 for hosts that keep hpsi in torch tensors: every rank
    of `group` owns the divide_all column block of hpsi (shape (m, lda[, 2])) and receives the others'. Works on
    any torch.distributed backend (NCCL on GPUs, gloo on CPU)."""
    import torch.distributed as dist

    world, rank = dist.get_world_size(group), dist.get_rank(group)
    for r in range(world):
        s, e = divide_all(m, world, r)
        if e > s:
            src = dist.get_global_rank(group, r) if group is not None else r
            dist.broadcast(hpsi[s:e], src=src, group=group)
    return hpsi


_CALLER_STREAMS = set()  # stream handles contexts were created on (HPsi(stream=...))


def _ptr(a, sync=True):
    """address of a numpy array / torch tensor / None / int"""
    if a is None:
        return None
    if isinstance(a, int):
        return ctypes.c_void_p(a)
    if isinstance(a, np.ndarray):
        if not a.flags.c_contiguous:
            raise QebError("array must be C-contiguous")
        return ctypes.c_void_p(a.ctypes.data)
    if hasattr(a, "data_ptr"):
        if not a.is_contiguous():
            raise QebError("tensor must be contiguous")
        if sync and getattr(a, "is_cuda", False):
            # The library's own streams are non-blocking: work the caller queued on torch's current stream (the fill or
            # the upload of this very tensor) is ordered before the library call here, as a Fortran caller's
            # synchronous host code is. A context created ON the caller's stream needs no such wait.
            import torch

            cs = torch.cuda.current_stream(a.device)
            if cs.cuda_stream not in _CALLER_STREAMS:
                cs.synchronize()
        return ctypes.c_void_p(a.data_ptr())
    raise QebError(f"unsupported array type {type(a)}")


def _i32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.int32)


def h_psi_multi(ctxs, lda, n, m, psi, hpsi):
    """qeb_h_psi_multi: one h_psi per context (k-point), all in flight together; lda/n/m ints or per-k lists; psi/hpsi lists
    of device tensors (or host arrays)."""
    nk = len(ctxs)
    if nk == 0:
        return hpsi
    ilist = lambda v: (ctypes.c_int * nk)(*([v] * nk if isinstance(v, int) else list(v)))  # noqa: E731
    hs = (ctypes.c_void_p * nk)(*[c._h.value for c in ctxs])
    _ptr(psi[0])  # orders the caller's torch stream before the call, once for all k-points
    pp = (ctypes.c_void_p * nk)(*[_ptr(a, sync=False).value for a in psi])
    ph = (ctypes.c_void_p * nk)(*[_ptr(a, sync=False).value for a in hpsi])
    L = ctxs[0]._L
    rc = L.qeb_h_psi_multi(nk, hs, ilist(lda), ilist(n), ilist(m), pp, ph)
    ctxs[0]._ck(rc, "qeb_h_psi_multi")
    return hpsi


class HPsi:
    """One context per (process, device): the explicit form of the module state
    h_psi reads in the reference."""

    def __init__(self, device: int = 0, stream: int | None = None, batch: int | None = None):
        self._L = load_library()
        self._h = ctypes.c_void_p()
        rc = self._L.qeb_create(ctypes.byref(self._h), device, ctypes.c_void_p(stream) if stream else None)
        if rc:
            self._h = None
            raise QebError(f"qeb_create failed (code {rc}): a CUDA device is required, there is no CPU fallback")
        if stream:
            _CALLER_STREAMS.add(int(stream))
        self.gamma_only = False
        self._keep = {}
        if batch:
            self.set_batch(batch)

    def close(self):
        if getattr(self, "_h", None):
            self._L.qeb_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc, who):
        if rc:
            raise QebError(f"{who} failed (code {rc}): {self._L.qeb_last_error(self._h).decode()}")

    # -- state setters ------------------------------------------------------
    def set_batch(self, nboxes: int):
        self._ck(self._L.qeb_set_batch(self._h, nboxes), "qeb_set_batch")

    def set_pipeline(self, chunk_cols: int = 0, min_cols: int = 128):
        self._ck(self._L.qeb_set_pipeline(self._h, chunk_cols, min_cols), "qeb_set_pipeline")

    def pin_host(self, arr):
        """Page-locks a host array for the host pipeline (cudaHostRegister); the caller unpins it before freeing."""
        self._ck(self._L.qeb_pin_host(self._h, _ptr(arr), arr.nbytes), "qeb_pin_host")

    def unpin_host(self, arr):
        self._ck(self._L.qeb_unpin_host(self._h, _ptr(arr)), "qeb_unpin_host")

    def set_async(self, flag: bool):
        self._ck(self._L.qeb_set_async(self._h, int(flag)), "qeb_set_async")

    def set_grid(self, nr1, nr2, nr3, nl, nlm=None, gamma_only=False, nr1x=None, nr2x=None, nr3x=None):
        nl = _i32(nl)
        nlm = _i32(nlm)
        self.gamma_only = bool(gamma_only)
        self.nr = (nr1, nr2, nr3)
        self._ck(self._L.qeb_set_grid(self._h, nr1, nr2, nr3, nr1x or nr1, nr2x or nr2, nr3x or nr3, nl.size, _ptr(nl),
                                      _ptr(nlm), int(gamma_only)), "qeb_set_grid")

    def set_grid_dist(self, nr1, nr2, nr3, nproc, mype, nr3p, nsw, iss, ismap, nl, nlm=None, gamma_only=False):
        """Per-rank descriptor of a slab-distributed dffts (fft_types.f90:35-185): see qeb_set_grid_dist."""
        nr3p, nsw, iss, ismap, nl, nlm = map(_i32, (nr3p, nsw, iss, ismap, nl, nlm))
        self.gamma_only = bool(gamma_only)
        self.nr = (nr1, nr2, nr3)
        self._ck(self._L.qeb_set_grid_dist(self._h, nr1, nr2, nr3, nproc, mype, _ptr(nr3p), _ptr(nsw), _ptr(iss),
                                           _ptr(ismap), ismap.size, nl.size, _ptr(nl), _ptr(nlm), int(gamma_only)),
                 "qeb_set_grid_dist")

    def nccl_init(self, role=COMM_RG, group=None):
        """Creates the library's own NCCL communicator for `role` over a torch.distributed group: rank 0 makes the
        unique id, torch.distributed broadcasts it (the Fortran host would MPI_Bcast it), every rank joins."""
        import torch
        import torch.distributed as dist

        rank, world = dist.get_rank(group), dist.get_world_size(group)
        buf = (ctypes.c_char * 128)()
        if rank == 0 and self._L.qeb_nccl_unique_id(buf):
            raise QebError("qeb_nccl_unique_id failed (libnccl not found? set QEB_NCCL_LIB)")
        t = torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8).clone()
        on_gpu = dist.get_backend(group) == "nccl"
        if on_gpu:
            t = t.cuda()
        src = dist.get_global_rank(group, 0) if group is not None else 0
        dist.broadcast(t, src=src, group=group)
        raw = bytes(t.cpu().numpy().tobytes())
        self._ck(self._L.qeb_nccl_init(self._h, role, ctypes.c_char_p(raw), world, rank), "qeb_nccl_init")

    def p2p_setup(self, group=None):
        """Peer-to-peer R&G exchange: export this rank's buffer handles, all-gather them over torch.distributed (a
        Fortran host would MPI_Allgather), import the peers', barrier. Call after set_grid_dist + set_kpoint."""
        import torch
        import torch.distributed as dist

        world = dist.get_world_size(group)
        blob = (ctypes.c_char * 256)()
        self._ck(self._L.qeb_p2p_export(self._h, blob), "qeb_p2p_export")
        mine = torch.frombuffer(bytearray(blob.raw), dtype=torch.uint8).clone()
        on_gpu = dist.get_backend(group) == "nccl"
        if on_gpu:
            mine = mine.cuda()
        allb = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allb, mine, group=group)
        raw = b"".join(bytes(t.cpu().numpy().tobytes()) for t in allb)
        self._ck(self._L.qeb_p2p_import(self._h, ctypes.c_char_p(raw), world), "qeb_p2p_import")
        dist.barrier(group=group)

    def set_kpoint(self, npw, npwx, igk, g2kin, vkb=None, nkb=0, gstart=1):
        igk = _i32(igk)
        if isinstance(g2kin, np.ndarray):
            g2kin = np.ascontiguousarray(g2kin, dtype=np.float64)
        self.npw, self.npwx, self.nkb = npw, npwx, nkb
        self._keep["vkb"] = vkb  # a device vkb is borrowed by the library
        self._ck(self._L.qeb_set_kpoint(self._h, npw, npwx, _ptr(igk), _ptr(g2kin), _ptr(vkb), nkb, gstart),
                 "qeb_set_kpoint")

    def set_vloc(self, vrs):
        self._ck(self._L.qeb_set_vloc(self._h, _ptr(vrs)), "qeb_set_vloc")

    def set_nonlocal(self, nh_of_atom, ofsbeta, deeq):
        """deeq: array laid out like Fortran deeq(nhm,nhm,nat) for the current spin, i.e. numpy shape (nat,nhm,nhm)."""
        nh = _i32(nh_of_atom)
        ofs = _i32(ofsbeta)
        deeq = np.ascontiguousarray(deeq, dtype=np.float64)
        nat = nh.size
        nhm = deeq.shape[-1] if deeq.size else 0
        self._ck(self._L.qeb_set_nonlocal(self._h, nat, nhm, _ptr(nh), _ptr(ofs), _ptr(deeq)), "qeb_set_nonlocal")

    def set_qq(self, qq_at):
        """qq_at: Fortran qq_at(nhm,nhm,nat), i.e. numpy shape (nat,nhm,nhm) (upflib/uspp.f90); after set_nonlocal."""
        qq = np.ascontiguousarray(qq_at, dtype=np.float64)
        self._ck(self._L.qeb_set_qq(self._h, qq.shape[0], qq.shape[-1] if qq.size else 0, _ptr(qq)), "qeb_set_qq")

    # -- device-side producers of the per-k-point state -------
    def set_gvectors(self, g, mill=None):
        """gvect::g (ngm,3) [= Fortran g(3,ngm)] in 2pi/alat and mill (ngm,3)."""
        g = np.ascontiguousarray(g, dtype=np.float64)
        self._ck(self._L.qeb_set_gvectors(self._h, g.shape[0], _ptr(g), _ptr(_i32(mill))), "qeb_set_gvectors")

    def g2_kin(self, xk, tpiba2, qcutz=0.0, ecfixed=0.0, q2sigma=1.0, out=None):
        """g2_kin(ik) (PW/src/g2_kin.f90:10): fills the context's g2kin on the device; returns it when out is given."""
        xk = np.ascontiguousarray(xk, dtype=np.float64)
        self._ck(self._L.qeb_g2_kin(self._h, _ptr(xk), tpiba2, qcutz, ecfixed, q2sigma, _ptr(out)), "qeb_g2_kin")
        return out

    def set_beta_tables(self, us):
        """us: dict as oracle.uspp_ref.synthetic_uspp (nh, nhm, nbetam, lmaxkb, nhtol/nhtolm/indv (nsp,nhm), nqx, dq,
        tab_beta (nsp,nbetam,nqx)) - numpy shapes are the Fortran arrays' memory order."""
        nh = _i32(us["nh"])
        a, b, c = _i32(us["nhtol"]), _i32(us["nhtolm"]), _i32(us["indv"])
        tab = np.ascontiguousarray(us["tab_beta"], dtype=np.float64)
        self._ck(self._L.qeb_set_beta_tables(self._h, nh.size, int(us["nhm"]), int(us["nbetam"]), int(us["lmaxkb"]),
                                             _ptr(nh), _ptr(a), _ptr(b), _ptr(c), int(us["nqx"]), float(us["dq"]),
                                             _ptr(tab)), "qeb_set_beta_tables")

    def set_structure(self, ityp, tau, nr, eigts):
        """ityp (nat,) 1-based, tau (nat,3) alat units, nr = dense grid dims, eigts = three (nat, 2nr_i+1) arrays."""
        ityp = _i32(ityp)
        tau = np.ascontiguousarray(tau, dtype=np.float64)
        e = [np.ascontiguousarray(x, dtype=np.complex128) for x in eigts]
        self._ck(self._L.qeb_set_structure(self._h, ityp.size, _ptr(ityp), _ptr(tau), nr[0], nr[1], nr[2], _ptr(e[0]),
                                           _ptr(e[1]), _ptr(e[2])), "qeb_set_structure")

    def init_us_2(self, q, tpiba, out=None):
        """init_us_2 (upflib/init_us_2_acc.f90:10): builds the context's vkb on the device; out (nkb,npwx) optional."""
        q = np.ascontiguousarray(q, dtype=np.float64)
        self._ck(self._L.qeb_init_us_2(self._h, _ptr(q), tpiba, _ptr(out)), "qeb_init_us_2")
        return out

    def sum_band_rho(self, lda, n, m, evc, w, rho):
        """density part of sum_band (PW/src/sum_band.f90:326-647): rho += sum_b w[b] |psi_b(r)|^2, w = wg/omega."""
        if isinstance(w, np.ndarray):
            w = np.ascontiguousarray(w, dtype=np.float64)
        self._ck(self._L.qeb_sum_band_rho(self._h, lda, n, m, _ptr(evc), _ptr(w), _ptr(rho)), "qeb_sum_band_rho")
        return rho

    def sum_band_rho_nc(self, lda, n, m, evc, w, rho, domag=True):
        """noncollinear branch of the sum_band loop (PW/src/sum_band.f90:555-575, get_rho_domag :699-729): evc (m, 2*lda)
        spinors; rho (4, nnr) when domag (density, m_x, m_y, m_z), (nnr,) otherwise."""
        if isinstance(w, np.ndarray):
            w = np.ascontiguousarray(w, dtype=np.float64)
        self._ck(self._L.qeb_sum_band_rho_nc(self._h, lda, n, m, _ptr(evc), _ptr(w), _ptr(rho), 1 if domag else 0),
                 "qeb_sum_band_rho_nc")
        return rho

    # -- block-Davidson building blocks; big arrays are torch CUDA tensors ---------
    def set_g_psi_diag(self, h_diag, s_diag):
        h_diag = np.ascontiguousarray(h_diag, dtype=np.float64)
        s_diag = np.ascontiguousarray(s_diag, dtype=np.float64)
        self._ck(self._L.qeb_set_g_psi_diag(self._h, h_diag.size, _ptr(h_diag), _ptr(s_diag)), "qeb_set_g_psi_diag")

    def g_psi(self, lda, n, m, npol, psi, e):
        """g_psi(lda,n,m,npol,psi,e) (PW/src/g_psi.f90:10): psi (device) is preconditioned in place."""
        e = np.ascontiguousarray(e, dtype=np.float64) if isinstance(e, np.ndarray) else e
        self._ck(self._L.qeb_g_psi(self._h, lda, n, m, npol, _ptr(psi), _ptr(e)), "qeb_g_psi")
        return psi

    def col_dot(self, n, m, a, lda, b, ldb):
        out = np.zeros(m)
        self._ck(self._L.qeb_col_dot(self._h, n, m, _ptr(a), lda, _ptr(b), ldb, _ptr(out)), "qeb_col_dot")
        return out

    def dav_gemm_cn(self, n, A, lda, ka, B, ldb, mb, C, ldc):
        self._ck(self._L.qeb_dav_gemm_cn(self._h, n, _ptr(A), lda, ka, _ptr(B), ldb, mb, _ptr(C), ldc), "qeb_dav_gemm_cn")
        return C

    def dav_gemm_nn(self, n, A, lda, k, B, ldb, m, C, ldc, accumulate=False):
        self._ck(self._L.qeb_dav_gemm_nn(self._h, n, _ptr(A), lda, k, _ptr(B), ldb, m, _ptr(C), ldc, int(accumulate)),
                 "qeb_dav_gemm_nn")
        return C

    def cegterg(self, npw, npwx, nvec, nvecx, evc, ethr, uspp, btype, lrot=False, npol=1):
        """cegterg / regterg (KS_Solvers/Davidson/cegterg.f90:18, regterg.f90:18) with the context's operators: evc
        (nvec, npwx) complex, host array or device tensor, is refined in place; returns (e, notcnv, dav_iter, nhpsi)."""
        e = np.zeros(nvec)
        bt = np.ascontiguousarray(btype, dtype=np.int32)
        out = (ctypes.c_int * 3)()
        pn, pi, ph = (ctypes.cast(ctypes.byref(out, 4 * k), ctypes.c_void_p) for k in range(3))
        if self.gamma_only:
            rc = self._L.qeb_regterg(self._h, npw, npwx, nvec, nvecx, _ptr(evc), float(ethr), int(bool(uspp)), _ptr(e), _ptr(bt),
                                     pn, int(bool(lrot)), pi, ph)
            self._ck(rc, "qeb_regterg")
        else:
            rc = self._L.qeb_cegterg(self._h, npw, npwx, nvec, nvecx, npol, _ptr(evc), float(ethr), int(bool(uspp)), _ptr(e),
                                     _ptr(bt), pn, int(bool(lrot)), pi, ph)
            self._ck(rc, "qeb_cegterg")
        return e, int(out[0]), int(out[1]), int(out[2])

    regterg = cegterg

    # -- reference entry points --------------------------------------------
    def s_psi(self, lda, n, m, psi, spsi):
        """s_psi(lda, n, m, psi, spsi) (PW/src/s_psi.f90:10): uses the becp the preceding h_psi/calbec left."""
        self._ck(self._L.qeb_s_psi(self._h, lda, n, m, _ptr(psi), _ptr(spsi)), "qeb_s_psi")
        return spsi

    def h_psi(self, lda, n, m, psi, hpsi):
        self._ck(self._L.qeb_h_psi(self._h, lda, n, m, _ptr(psi), _ptr(hpsi)), "qeb_h_psi")
        return hpsi

    def hs_psi(self, lda, n, m, psi, hpsi, spsi):
        """hs_psi(lda,n,m,psi,hpsi,spsi) (PW/src/hs_psi.f90:7); m = 1 is hs_1psi (PW/src/hs_1psi.f90:7)"""
        self._ck(self._L.qeb_hs_psi(self._h, lda, n, m, _ptr(psi), _ptr(hpsi), _ptr(spsi)), "qeb_hs_psi")
        return hpsi, spsi

    def h_psi_bgrp(self, lda, n, m, psi, hpsi, nbgrp, my_bgrp):
        self._ck(self._L.qeb_h_psi_bgrp(self._h, lda, n, m, _ptr(psi), _ptr(hpsi), nbgrp, my_bgrp), "qeb_h_psi_bgrp")
        return hpsi

    def vloc_psi_k(self, lda, n, m, psi, v, hpsi):
        self._ck(self._L.qeb_vloc_psi_k(self._h, lda, n, m, _ptr(psi), _ptr(v), _ptr(hpsi)), "qeb_vloc_psi_k")
        return hpsi

    def vloc_psi_gamma(self, lda, n, m, psi, v, hpsi):
        self._ck(self._L.qeb_vloc_psi_gamma(self._h, lda, n, m, _ptr(psi), _ptr(v), _ptr(hpsi)), "qeb_vloc_psi_gamma")
        return hpsi

    def set_vloc_nc(self, v4):
        """the 2x2 potential v(dffts%nnr, 4) of the noncollinear magnetic case (vloc_psi_acc.f90:291)"""
        self._ck(self._L.qeb_set_vloc_nc(self._h, _ptr(v4)), "qeb_set_vloc_nc")

    def vloc_psi_nc(self, lda, n, m, psi, v, hpsi):
        """vloc_psi_nc(lda, n, m, psi, v, hpsi), domag branch: psi(lda*2, m), hpsi(lda, 2, m) INOUT, v(nnr, 4) or None"""
        self._ck(self._L.qeb_vloc_psi_nc(self._h, lda, n, m, _ptr(psi), _ptr(v), _ptr(hpsi)), "qeb_vloc_psi_nc")
        return hpsi

    def set_nonlocal_nc(self, deeq_nc):
        """deeq_nc(nhm,nhm,nat,4) complex (memory (4, nat, nhm, nhm)); nat/nhm/nh/ofsbeta from set_nonlocal"""
        self._ck(self._L.qeb_set_nonlocal_nc(self._h, _ptr(deeq_nc)), "qeb_set_nonlocal_nc")

    def h_psi_nc(self, lda, n, m, psi, hpsi):
        """h_psi for noncollinear spinors with magnetisation: psi(lda*2, m) -> hpsi(lda*2, m)"""
        self._ck(self._L.qeb_h_psi_nc(self._h, lda, n, m, _ptr(psi), _ptr(hpsi)), "qeb_h_psi_nc")
        return hpsi

    def set_qq_nc(self, qq_nc):
        """per-atom 2x2-spin overlap (nhm,nhm,nat,4) complex (memory (4, nat, nhm, nhm)) of s_psi_nc"""
        self._ck(self._L.qeb_set_qq_nc(self._h, _ptr(qq_nc)), "qeb_set_qq_nc")

    def s_psi_nc(self, lda, n, m, psi, spsi):
        """s_psi for spinors: psi(lda*2, m) -> spsi(lda*2, m), becp%nc from the preceding h_psi_nc"""
        self._ck(self._L.qeb_s_psi_nc(self._h, lda, n, m, _ptr(psi), _ptr(spsi)), "qeb_s_psi_nc")
        return spsi

    def invfft(self, kind, f, howmany=1):
        """invfft(kind, f, dfft, howmany): G->R, in place."""
        self._ck(self._L.qeb_fft(self._h, +1, FFT_WAVE if kind in ("Wave", FFT_WAVE) else FFT_RHO, _ptr(f), howmany),
                 "qeb_fft")
        return f

    def fwfft(self, kind, f, howmany=1):
        """fwfft(kind, f, dfft, howmany): R->G, scaled, in place."""
        self._ck(self._L.qeb_fft(self._h, -1, FFT_WAVE if kind in ("Wave", FFT_WAVE) else FFT_RHO, _ptr(f), howmany),
                 "qeb_fft")
        return f

    def calbec(self, npw, beta, psi, betapsi, nbnd, lda=None, nkb=None):
        """calbec(npw, beta, psi, betapsi, nbnd). beta None -> the context's vkb."""
        if lda is None:
            lda = psi.shape[-1]
        if nkb is None:
            nkb = self.nkb if beta is None else beta.shape[0]
        self._ck(self._L.qeb_calbec(self._h, npw, _ptr(beta), nkb, _ptr(psi), lda, nbnd, _ptr(betapsi)), "qeb_calbec")
        return betapsi

    def calbec_static(self, npw, beta, psi, betapsi, nbnd, lda=None, nkb=None):
        """qeb_calbec_static: calbec on device arrays with the kernel's scratch in the library's static device array;
        allocates nothing (no reference counterpart; see include/qeb200.h)."""
        if lda is None:
            lda = psi.shape[-1]
        if nkb is None:
            nkb = self.nkb if beta is None else beta.shape[0]
        self._ck(self._L.qeb_calbec_static(self._h, npw, _ptr(beta), nkb, _ptr(psi), lda, nbnd, _ptr(betapsi)),
                 "qeb_calbec_static")
        return betapsi

    def add_vuspsi(self, lda, n, m, hpsi, becp=None):
        self._ck(self._L.qeb_add_vuspsi(self._h, lda, n, m, _ptr(hpsi), _ptr(becp)), "qeb_add_vuspsi")
        return hpsi

    def get_becp(self, m):
        out = np.empty((m, self.nkb), dtype=np.float64 if self.gamma_only else np.complex128)
        self._ck(self._L.qeb_get_becp(self._h, _ptr(out), m), "qeb_get_becp")
        return out

    def nonlocal_only(self, lda, n, m, psi, hpsi):
        self._ck(self._L.qeb_nonlocal_only(self._h, lda, n, m, _ptr(psi), _ptr(hpsi)), "qeb_nonlocal_only")

    def vloc_kernels_per_batch(self):
        return int(self._L.qeb_vloc_kernels_per_batch(self._h))

    def stats(self):
        vb, gf = ctypes.c_double(), ctypes.c_double()
        nst, nxw, nb = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        self._ck(self._L.qeb_get_stats(self._h, ctypes.byref(vb), ctypes.byref(gf), ctypes.byref(nst),
                                       ctypes.byref(nxw), ctypes.byref(nb)), "qeb_get_stats")
        return dict(vloc_bytes_per_band=vb.value, gemm_flops_per_band=gf.value, nst=nst.value, nxw=nxw.value,
                    batch=nb.value)

    # -- convenience: load everything from an oracle-style descriptor --------
    def setup_from(self, d, inp, lda=None):
        """d: object with nr1..3, nl, nlm, gamma_only, npw, igk, g2kin, gstart (1-based maps as in the reference);
        inp: dict with vrs, vkb (nkb, npwx), deeq, nh_at, ofsbeta."""
        self.set_grid(d.nr1, d.nr2, d.nr3, d.nl, d.nlm if d.gamma_only else None, d.gamma_only)
        npwx = lda if lda is not None else inp["lda"]
        vkb = inp.get("vkb")
        nkb = 0 if vkb is None else vkb.shape[0]
        self.set_kpoint(d.npw, npwx, None if d.gamma_only else d.igk, d.g2kin, vkb if nkb else None, nkb, d.gstart)
        self.set_vloc(inp["vrs"])
        if nkb:
            self.set_nonlocal(inp["nh_at"], inp["ofsbeta"], inp["deeq"])
