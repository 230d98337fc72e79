"""ORACLE (test infrastructure, not product code).

CPU restatement of how the reference builds everything `h_psi` reads from
module state: FFT grid dimensions, the |G|-sorted G-vector list, the sphere->box
index maps nl/nlm, the per-k list igk + g2kin, and the wave stick/plane masks
(isind/iplw) used by the "reduced" 3-D FFT.

Follows (paths relative to /root/reference):
  realspace_grid_init / grid_set   FFTXlib/src/fft_types.f90:1070-1197
  good_fft_order / allowed         FFTXlib/src/fft_support.f90:61-165
  ggen (loop order, eps8 snap)     Modules/recvec_subs.f90:111-258
  ggens (smooth prefix)            Modules/recvec_subs.f90:263-311
  hpsort_eps                       Modules/sort.f90:9-134   (C: oracle/c/qe_sort.c)
  fft_set_nl                       FFTXlib/src/fft_ggen.f90:28-108
  gk_sort                          PW/src/gk_sort.f90:9-98
  g2_kin                           PW/src/g2_kin.f90:37-43
  sticks_map_set / get_sticks      FFTXlib/src/stick_base.f90:179-262,550-601
  serial isind/iplw                FFTXlib/src/fft_types.f90:814-851
  gkcut                            PW/src/data_structure.f90:48-68
  recips                           FFTXlib/tests/recips.f90

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import
this module. All index arrays returned here are 1-based (Fortran values), as the
C ABI expects them.
"""
from __future__ import annotations

import ctypes
import os
from dataclasses import dataclass, field

import numpy as np

EPS8 = 1.0e-8
_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def _lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            import subprocess

            subprocess.check_call(["make", "-C", _HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
        _LIB = ctypes.CDLL(path)
        _LIB.qeo_hpsort_eps.argtypes = [
            ctypes.c_int,
            ctypes.POINTER(ctypes.c_double),
            ctypes.POINTER(ctypes.c_int),
            ctypes.c_double,
        ]
        _LIB.qeo_hpsort_eps.restype = None
    return _LIB


def hpsort_eps(ra: np.ndarray, ind: np.ndarray | None = None, eps: float = EPS8):
    """Returns (sorted ra, carried 1-based index array). Modules/sort.f90:9."""
    ra = np.ascontiguousarray(ra, dtype=np.float64).copy()
    n = ra.size
    if ind is None:
        ind = np.zeros(max(n, 1), dtype=np.int32)
    else:
        ind = np.ascontiguousarray(ind, dtype=np.int32).copy()
    if n > 0:
        _lib().qeo_hpsort_eps(
            n,
            ra.ctypes.data_as(ctypes.POINTER(ctypes.c_double)),
            ind.ctypes.data_as(ctypes.POINTER(ctypes.c_int)),
            eps,
        )
    return ra, ind[:n]


def allowed(nr: int) -> bool:
    """fft_support.f90:61-120, non-ESSL branch: only factors 2,3,5."""
    mr = nr
    for fac in (2, 3, 5, 7, 11):
        pw = 0
        while mr % fac == 0 and mr > 1:
            mr //= fac
            pw += 1
        if fac in (7, 11) and pw > 0:
            return False
    return mr == 1


def good_fft_order(nr: int, np_: int | None = None) -> int:
    """fft_support.f90:124-160."""
    new = nr
    while (not allowed(new)) or (np_ is not None and new % np_ != 0):
        new += 1
        if new > 16385:
            raise ValueError("fft order too large")
    return new


def recips(at: np.ndarray) -> np.ndarray:
    """bg(:,i) in units of 2pi/alat with at(:,i) in units of alat (columns)."""
    a1, a2, a3 = at[:, 0], at[:, 1], at[:, 2]
    den = np.dot(a1, np.cross(a2, a3))
    b1 = np.cross(a2, a3) / den
    b2 = np.cross(a3, a1) / den
    b3 = np.cross(a1, a2) / den
    return np.stack([b1, b2, b3], axis=1)


@dataclass
class Descriptor:
    """Everything the reference keeps in dffts + gvect + wvfct for one k-point."""

    at: np.ndarray
    bg: np.ndarray
    alat: float
    tpiba2: float
    ecutwfc: float
    gamma_only: bool
    nr1: int
    nr2: int
    nr3: int
    ngm: int  # G-vectors of the smooth grid (half sphere if gamma_only)
    mill: np.ndarray  # (ngm,3) Miller indices
    g: np.ndarray  # (ngm,3) in 2pi/alat
    gg: np.ndarray  # (ngm,)
    gstart: int  # 2 if G(1)=0 else 1
    nl: np.ndarray  # (ngm,) 1-based box index
    nlm: np.ndarray | None  # (ngm,) 1-based, gamma only
    xk: np.ndarray  # k in 2pi/alat
    npw: int
    igk: np.ndarray  # (npw,) 1-based into G list
    g2kin: np.ndarray  # (npw,) Ry
    isind: np.ndarray  # (nr1*nr2,) 0/1 wave-stick mask (box column index, x fastest)
    iplw: np.ndarray  # (nr1,) 0/1 active x-planes
    nst: int = 0
    extra: dict = field(default_factory=dict)

    @property
    def nnr(self) -> int:
        return self.nr1 * self.nr2 * self.nr3


def _grid_dims(at, bg, gcut):
    """realspace_grid_init + grid_set (fft_types.f90:1086-1122,1128-1197)."""
    est = [int(np.sqrt(gcut) * np.sqrt(np.sum(at[:, i] ** 2))) + 1 for i in range(3)]
    i = np.arange(-est[0], est[0] + 1)
    j = np.arange(-est[1], est[1] + 1)
    k = np.arange(-est[2], est[2] + 1)
    I, J, K = np.meshgrid(i, j, k, indexing="ij")
    gx = I * bg[0, 0] + J * bg[0, 1] + K * bg[0, 2]
    gy = I * bg[1, 0] + J * bg[1, 1] + K * bg[1, 2]
    gz = I * bg[2, 0] + J * bg[2, 1] + K * bg[2, 2]
    gsq = gx**2 + gy**2 + gz**2
    m = gsq < gcut  # strict, fft_types.f90:1171
    nb = [int(np.abs(I[m]).max()), int(np.abs(J[m]).max()), int(np.abs(K[m]).max())]
    return tuple(good_fft_order(2 * b + 1) for b in nb), nb


def _gen_g(nr, bg, gcut, gamma_only):
    """ggen loop nest (recvec_subs.f90:113-190) restricted to |G|^2<=gcut."""
    ni, nj, nk = (nr[0] - 1) // 2, (nr[1] - 1) // 2, (nr[2] - 1) // 2
    i = np.arange(0 if gamma_only else -ni, ni + 1)
    j = np.arange(-nj, nj + 1)
    k = np.arange(-nk, nk + 1)
    I, J, K = np.meshgrid(i, j, k, indexing="ij")  # C order == i outer, k inner
    # t = (i*bg1 + j*bg2) + k*bg3 with the reference's association
    t = [(I * bg[c, 0] + J * bg[c, 1]) + K * bg[c, 2] for c in range(3)]
    tt = t[0] ** 2 + t[1] ** 2 + t[2] ** 2
    keep = tt <= gcut
    if gamma_only:
        keep &= ~((I == 0) & (J < 0))
        keep &= ~((I == 0) & (J == 0) & (K < 0))
    I, J, K, tt = I[keep], J[keep], K[keep], tt[keep]
    g2sort = np.where(tt > EPS8, tt, 0.0)
    _, igsrt = hpsort_eps(g2sort)
    o = igsrt - 1
    mill = np.stack([I[o], J[o], K[o]], axis=1).astype(np.int32)
    g = mill[:, 0:1] * bg[:, 0][None, :] + mill[:, 1:2] * bg[:, 1][None, :] + mill[:, 2:3] * bg[:, 2][None, :]
    gg = np.sum(g**2, axis=1)
    return mill, g, gg


def _stick_map(nr, bg, gcut, gamma_only):
    """st(i1,i2) > 0 where the column holds a G with |G|^2<=gcut; Gamma: the
    half-space rule then mirrored to (-i1,-i2) (stick_base.f90:226-245,594)."""
    ub = [(nr[0] - 1) // 2, (nr[1] - 1) // 2, (nr[2] - 1) // 2]
    i = np.arange(-ub[0], ub[0] + 1)
    j = np.arange(-ub[1], ub[1] + 1)
    k = np.arange(-ub[2], ub[2] + 1)
    I, J, K = np.meshgrid(i, j, k, indexing="ij")
    amod = (
        (I * bg[0, 0] + J * bg[0, 1] + K * bg[0, 2]) ** 2
        + (I * bg[1, 0] + J * bg[1, 1] + K * bg[1, 2]) ** 2
        + (I * bg[2, 0] + J * bg[2, 1] + K * bg[2, 2]) ** 2
    )
    inside = amod <= gcut
    if gamma_only:
        inside &= I >= 0
        inside &= ~((I == 0) & (J < 0))
        inside &= ~((I == 0) & (J == 0) & (K < 0))
    st = inside.sum(axis=2)  # (2ub1+1, 2ub2+1) counts
    ngm = int(inside.sum())
    has = st > 0
    if gamma_only:
        has = has | has[::-1, ::-1]
    return has, ub, ngm


def build_descriptor(
    at_cols,
    alat: float,
    ecutwfc: float,
    xk=(0.0, 0.0, 0.0),
    gamma_only: bool = False,
    ecutrho: float | None = None,
    kmax: float | None = None,
) -> Descriptor:
    """at_cols: (3,3) array whose COLUMNS are a_i in units of alat.
    xk in units of 2pi/alat (cartesian). ecutrho only matters for deciding the
    smooth cutoff: gcutms = min(ecutrho, 4*ecutwfc)/tpiba2 (data_structure /
    gvecs_init); h_psi only touches the smooth grid."""
    at = np.asarray(at_cols, dtype=np.float64)
    bg = recips(at)
    tpiba = 2.0 * np.pi / alat
    tpiba2 = tpiba * tpiba
    gcutw = ecutwfc / tpiba2
    gcutms = 4.0 * gcutw
    if ecutrho is not None and ecutrho < 4.0 * ecutwfc:
        gcutms = ecutrho / tpiba2
    xk = np.asarray(xk, dtype=np.float64)
    knorm = float(np.sqrt(np.sum(xk**2)))
    if gamma_only and knorm > 0:
        raise ValueError("gamma_only requires k=0")
    if kmax is None:
        kmax = knorm
    gkcut = (np.sqrt(gcutw) + kmax) ** 2  # data_structure.f90:59-64

    (nr1, nr2, nr3), _nb = _grid_dims(at, bg, gcutms)
    nr = (nr1, nr2, nr3)
    mill, g, gg = _gen_g(nr, bg, gcutms, gamma_only)
    # ggens: prefix with gg <= gcutms (already restricted by tt<=gcut; re-check with gg)
    keep = gg <= gcutms
    if not keep.all():
        first_bad = int(np.argmin(keep))
        mill, g, gg = mill[:first_bad], g[:first_bad], gg[:first_bad]
    ngm = mill.shape[0]
    gstart = 2 if gg[0] <= EPS8 else 1

    # fft_set_nl (serial): nl = 1 + n1 + n2*nr1x + n3*nr1x*nr2x
    n1 = np.where(mill[:, 0] < 0, mill[:, 0] + nr1, mill[:, 0]).astype(np.int64)
    n2 = np.where(mill[:, 1] < 0, mill[:, 1] + nr2, mill[:, 1]).astype(np.int64)
    n3 = np.where(mill[:, 2] < 0, mill[:, 2] + nr3, mill[:, 2]).astype(np.int64)
    nl = (1 + n1 + n2 * nr1 + n3 * nr1 * nr2).astype(np.int32)
    nlm = None
    if gamma_only:
        m1 = np.where(-mill[:, 0] < 0, -mill[:, 0] + nr1, -mill[:, 0]).astype(np.int64)
        m2 = np.where(-mill[:, 1] < 0, -mill[:, 1] + nr2, -mill[:, 1]).astype(np.int64)
        m3 = np.where(-mill[:, 2] < 0, -mill[:, 2] + nr3, -mill[:, 2]).astype(np.int64)
        nlm = (1 + m1 + m2 * nr1 + m3 * nr1 * nr2).astype(np.int32)

    # gk_sort
    kg = g + xk[None, :]
    q = np.sum(kg**2, axis=1)
    q = np.where(q <= EPS8, 0.0, q)
    sel = np.nonzero(q <= gcutw)[0]
    gk = q[sel]
    igk = (sel + 1).astype(np.int32)
    if knorm**2 > EPS8:
        gk, igk = hpsort_eps(gk, igk)
    npw = igk.size
    kgs = g[igk - 1] + xk[None, :]
    g2kin = (kgs[:, 0] * kgs[:, 0] + kgs[:, 1] * kgs[:, 1] + kgs[:, 2] * kgs[:, 2]) * tpiba2

    # wave sticks (serial descriptor): isind over box columns, iplw over x
    has, ub, _ngw = _stick_map(nr, bg, gkcut, gamma_only)
    isind = np.zeros(nr1 * nr2, dtype=np.int32)
    iplw = np.zeros(nr1, dtype=np.int32)
    ii, jj = np.nonzero(has)
    i1 = ii - ub[0]
    i2 = jj - ub[1]
    m1 = np.where(i1 < 0, i1 + nr1, i1)
    m2 = np.where(i2 < 0, i2 + nr2, i2)
    isind[m1 + m2 * nr1] = 1
    iplw[m1] = 1
    nst = int(isind.sum())

    d = Descriptor(
        at=at, bg=bg, alat=alat, tpiba2=tpiba2, ecutwfc=ecutwfc, gamma_only=gamma_only,
        nr1=nr1, nr2=nr2, nr3=nr3, ngm=ngm, mill=mill, g=g, gg=gg, gstart=gstart,
        nl=nl, nlm=nlm, xk=xk, npw=npw, igk=igk, g2kin=g2kin, isind=isind, iplw=iplw, nst=nst,
    )
    d.extra["gcutw"] = gcutw
    d.extra["gcutms"] = gcutms
    d.extra["gkcut"] = gkcut
    d.extra["ngw_sticks"] = _ngw
    return d


# ----------------------------------------------------------------------------
# The five BASELINE.json configurations + small cases
# ----------------------------------------------------------------------------
def _fcc(a):
    # QE ibrav=2: a1=(-1,0,1)/2, a2=(0,1,1)/2, a3=(-1,1,0)/2 (columns)
    return np.array([[-0.5, 0.0, -0.5], [0.0, 0.5, 0.5], [0.5, 0.5, 0.0]])


def _ortho(a, b_over_a, c_over_a):
    return np.diag([1.0, b_over_a, c_over_a])


CONFIGS = {
    # name: dict(at, alat, ecutwfc, ecutrho, xk, gamma, nbnd, nat, nh)
    "si2": dict(at=_fcc(10.20), alat=10.20, ecutwfc=30.0, ecutrho=None, xk=(0.125, 0.125, 0.125),
                gamma=False, nbnd=4, nat=2, nh=4),
    "al4": dict(at=np.eye(3), alat=7.65, ecutwfc=30.0, ecutrho=None, xk=(0.0625, 0.0625, 0.0625),
                gamma=False, nbnd=10, nat=4, nh=4),
    "h2o64": dict(at=np.eye(3), alat=23.47, ecutwfc=80.0, ecutrho=None, xk=(0, 0, 0),
                  gamma=True, nbnd=256, nat=192, nh=None, nh_list=[8] * 64 + [2] * 128),
    "ausurf112": dict(at=_ortho(38.7583, 0.494393, 1.569966), alat=38.7583, ecutwfc=25.0, ecutrho=200.0,
                      xk=(0.25, 0.25, 0.0), gamma=False, nbnd=800, nat=112, nh=18),
    "si512": dict(at=np.eye(3), alat=40.8, ecutwfc=30.0, ecutrho=None, xk=(0, 0, 0),
                  gamma=True, nbnd=1024, nat=512, nh=4),
    # small cases for fast tests (test-suite-like Si ecut 12: 15^3, 180ish PWs)
    "si2_small": dict(at=_fcc(10.20), alat=10.20, ecutwfc=12.0, ecutrho=None, xk=(0.25, 0.25, 0.25),
                      gamma=False, nbnd=4, nat=2, nh=4),
    "si2_small_gamma": dict(at=_fcc(10.20), alat=10.20, ecutwfc=12.0, ecutrho=None, xk=(0, 0, 0),
                            gamma=True, nbnd=5, nat=2, nh=4),
    "ortho_small": dict(at=_ortho(9.0, 0.8, 1.3), alat=9.0, ecutwfc=14.0, ecutrho=None, xk=(0.1, 0.2, 0.05),
                        gamma=False, nbnd=7, nat=3, nh=8),
    "ortho_small_gamma": dict(at=_ortho(9.0, 0.8, 1.3), alat=9.0, ecutwfc=14.0, ecutrho=None, xk=(0, 0, 0),
                              gamma=True, nbnd=7, nat=3, nh=8),
}


def config_descriptor(name: str) -> Descriptor:
    c = CONFIGS[name]
    d = build_descriptor(c["at"], c["alat"], c["ecutwfc"], xk=c["xk"], gamma_only=c["gamma"], ecutrho=c["ecutrho"])
    d.extra["config"] = dict(c)
    d.extra["name"] = name
    return d
