"""Host-side mirror of the library's descriptor builder (include/qeb200.h, qeb_desc_*): what fft_type_init + ggen +
fft_set_nl + gk_sort + g2_kin give a PWscf run for one k-point (FFTXlib/src/fft_types.f90:947-1197,
Modules/recvec_subs.f90:111-258, FFTXlib/src/fft_ggen.f90:28-108, PW/src/gk_sort.f90, PW/src/g2_kin.f90:37-43) and
the slab distribution of the grid over ranks (FFTXlib/src/stick_base.f90:214-601, fft_types.f90:491-943).
Needs no GPU. Used by bench.py and by hosts that do not run the reference's own setup code."""
from __future__ import annotations

import ctypes
from types import SimpleNamespace

import numpy as np

from .hpsi import QebError, load_library

NL, NLM, IGK, G2KIN, MILL, GG, ISIND, IPLW, NR3P, NSW, NSP, ISS, ISMAP, IG_L2G, PW_L2G = range(15)

# The five BASELINE.json configurations: lattice (columns = a_i / alat), alat (bohr),
# cutoffs (Ry), first k-point (2pi/alat), bands, atoms and projectors per atom of the assumed pseudopotentials.
_FCC = [[-0.5, 0.0, -0.5], [0.0, 0.5, 0.5], [0.5, 0.5, 0.0]]
_CUBE = [[1.0, 0, 0], [0, 1.0, 0], [0, 0, 1.0]]
WORKLOADS = {
    "si2": dict(at=_FCC, alat=10.20, ecutwfc=30.0, ecutrho=0.0, xk=(0.125, 0.125, 0.125), gamma=False, nbnd=4,
                nh=[4] * 2),
    "al4": dict(at=_CUBE, alat=7.65, ecutwfc=30.0, ecutrho=0.0, xk=(0.0625, 0.0625, 0.0625), gamma=False, nbnd=10,
                nh=[4] * 4),
    "h2o64": dict(at=_CUBE, alat=23.47, ecutwfc=80.0, ecutrho=0.0, xk=(0, 0, 0), gamma=True, nbnd=256,
                  nh=[8] * 64 + [2] * 128),
    "ausurf112": dict(at=[[1.0, 0, 0], [0, 0.494393, 0], [0, 0, 1.569966]], alat=38.7583, ecutwfc=25.0, ecutrho=200.0,
                      xk=(0.25, 0.25, 0.0), gamma=False, nbnd=800, nh=[18] * 112),
    "si512": dict(at=_CUBE, alat=40.8, ecutwfc=30.0, ecutrho=0.0, xk=(0, 0, 0), gamma=True, nbnd=1024, nh=[4] * 512),
}


def good_fft_order(nr: int) -> int:
    return load_library().qeb_good_fft_order(int(nr))


class Descriptor:
    """One k-point's dffts + gvect + wvfct data. Attributes follow the reference's names (1-based maps)."""

    def __init__(self, at_cols, alat, ecutwfc, ecutrho=0.0, xk=(0.0, 0.0, 0.0), gamma_only=False):
        L = load_library()
        P, I, D = ctypes.c_void_p, ctypes.c_int, ctypes.c_double
        L.qeb_desc_create.argtypes = [ctypes.POINTER(P), P, D, D, D, P, I]
        L.qeb_desc_destroy.argtypes = [P]
        L.qeb_desc_error.argtypes = [P]
        L.qeb_desc_error.restype = ctypes.c_char_p
        L.qeb_desc_dims.argtypes = [P] + [ctypes.POINTER(I)] * 7
        L.qeb_desc_distribute.argtypes = [P, I]
        L.qeb_desc_rank_dims.argtypes = [P, I] + [ctypes.POINTER(I)] * 4
        L.qeb_desc_get.argtypes = [P, I, I, P]
        self._L = L
        self._h = P()
        at = np.ascontiguousarray(at_cols, dtype=np.float64).reshape(3, 3)
        k = np.ascontiguousarray(xk, dtype=np.float64)
        rc = L.qeb_desc_create(ctypes.byref(self._h), at.ctypes.data, float(alat), float(ecutwfc), float(ecutrho or 0.0),
                               k.ctypes.data, int(bool(gamma_only)))
        if rc:
            self._h = None
            raise QebError(f"qeb_desc_create failed (code {rc})")
        v = [I() for _ in range(7)]
        L.qeb_desc_dims(self._h, *[ctypes.byref(x) for x in v])
        self.nr1, self.nr2, self.nr3, self.ngm, self.npw, self.nst, self.gstart = [x.value for x in v]
        self.gamma_only = bool(gamma_only)
        self.nnr = self.nr1 * self.nr2 * self.nr3
        self.xk = k
        self.nl = self._get(NL, -1, self.ngm, np.int32)
        self.nlm = self._get(NLM, -1, self.ngm, np.int32) if gamma_only else None
        self.igk = self._get(IGK, -1, self.npw, np.int32)
        self.g2kin = self._get(G2KIN, -1, self.npw, np.float64)
        self.mill = self._get(MILL, -1, 3 * self.ngm, np.int32).reshape(self.ngm, 3)
        self.gg = self._get(GG, -1, self.ngm, np.float64)
        self.isind = self._get(ISIND, -1, self.nr1 * self.nr2, np.int32)
        self.iplw = self._get(IPLW, -1, self.nr1, np.int32)
        self.nproc = 0

    def _get(self, what, rank, n, dtype):
        out = np.zeros(max(int(n), 1), dtype=dtype)
        rc = self._L.qeb_desc_get(self._h, what, rank, out.ctypes.data)
        if rc:
            raise QebError(f"qeb_desc_get({what}, rank {rank}) failed (code {rc})")
        return out[:n]

    def distribute(self, nproc: int):
        """Slab (R&G) distribution over nproc ranks: fills nr3p, nsw, nsp, iss, ismap and ranks[r] with the rank's
        ngm, npw, gstart, nl, nlm, igk, g2kin, ig_l2g, pw_l2g."""
        rc = self._L.qeb_desc_distribute(self._h, int(nproc))
        if rc:
            raise QebError(f"qeb_desc_distribute failed: {self._L.qeb_desc_error(self._h).decode()}")
        self.nproc = nproc
        I = ctypes.c_int
        for name, what in (("nr3p", NR3P), ("nsw", NSW), ("nsp", NSP), ("iss", ISS)):
            setattr(self, name, self._get(what, -1, nproc, np.int32))
        self.i0r3p = np.concatenate([[0], np.cumsum(self.nr3p)[:-1]]).astype(np.int32)
        self.ranks = []
        for r in range(nproc):
            v = [I() for _ in range(4)]
            self._L.qeb_desc_rank_dims(self._h, r, *[ctypes.byref(x) for x in v])
            ngm, npw, gstart, nismap = [x.value for x in v]
            self.ranks.append(SimpleNamespace(
                mype=r, ngm=ngm, npw=npw, gstart=gstart,
                nl=self._get(NL, r, ngm, np.int32), nlm=self._get(NLM, r, ngm, np.int32) if self.gamma_only else None,
                igk=self._get(IGK, r, npw, np.int32), g2kin=self._get(G2KIN, r, npw, np.float64),
                ig_l2g=self._get(IG_L2G, r, ngm, np.int32), pw_l2g=self._get(PW_L2G, r, npw, np.int32)))
        self.ismap = self._get(ISMAP, -1, nismap, np.int32)
        return self

    def close(self):
        if getattr(self, "_h", None):
            self._L.qeb_desc_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def workload_descriptor(name: str, xk=None) -> tuple[Descriptor, dict]:
    w = WORKLOADS[name]
    d = Descriptor(w["at"], w["alat"], w["ecutwfc"], w["ecutrho"], w["xk"] if xk is None else xk, w["gamma"])
    return d, w
