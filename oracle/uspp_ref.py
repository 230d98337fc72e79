"""TEST INFRASTRUCTURE (oracle): CPU restatement of the per-k-point producers of h_psi's inputs
(SURVEY.md section 8 row f2). Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline may import this.

  g2_kin        PW/src/g2_kin.f90:10-62
  struc_fact    PW/src/struct_fact.f90:60-87 (the eigts1/2/3 phase tables only)
  ylmr2         upflib/ylmr2.f90:9-138 (CPU branch)
  interp_beta   upflib/beta_mod.f90:130-173
  init_us_2     upflib/init_us_2_acc.f90:10-172

Parity pinning: the reference has no fixture for these routines and cannot be built here (Fortran);
ylmr2 is cross-checked against scipy's complex spherical harmonics in tests/test_oracle.py."""
from __future__ import annotations

import math

import numpy as np

TPI = 2.0 * math.pi
FPI = 4.0 * math.pi


def g2_kin(g, igk, xk, tpiba2, qcutz=0.0, ecfixed=0.0, q2sigma=1.0):
    """g: (ngm,3) 2pi/alat; igk 1-based. g2_kin.f90:37-58."""
    kg = g[np.asarray(igk) - 1]
    a, b, c = xk[0] + kg[:, 0], xk[1] + kg[:, 1], xk[2] + kg[:, 2]
    g2 = (a * a + b * b + c * c) * tpiba2
    if qcutz > 0.0:
        erf = np.vectorize(math.erf)
        g2 = g2 + qcutz * (1.0 + erf((g2 - ecfixed) / q2sigma))
    return g2


def eigts(bg, tau, nr):
    """eigts_i(-nr_i:nr_i, na) = exp(-i 2pi n (bg_i . tau_na)); bg columns = b_i (as oracle.descriptor.recips),
    tau (nat,3) alat units. Returns three arrays of shape (nat, 2*nr_i+1) = Fortran eigts_i(-nr:nr, nat) memory.
    struct_fact.f90:68-87."""
    out = []
    for ipol in range(3):
        bgtau = tau @ bg[:, ipol]  # bg(1,ipol)*tau(1,na) + ...
        n = np.arange(-nr[ipol], nr[ipol] + 1)
        arg = TPI * n[None, :] * bgtau[:, None]
        out.append(np.cos(arg) - 1j * np.sin(arg))
    return out


def ylmr2(lmax2, g, gg):
    """g (ng,3), gg (ng,) -> ylm (ng, lmax2). ylmr2.f90:55-134."""
    ng = g.shape[0]
    lmax = int(round(math.sqrt(lmax2))) - 1
    assert (lmax + 1) ** 2 == lmax2
    ylm = np.zeros((ng, lmax2))
    if lmax == 0:
        ylm[:, 0] = math.sqrt(1.0 / FPI)
        return ylm
    eps = 1.0e-9
    gmod = np.sqrt(gg)
    cost = np.where(gmod < eps, 0.0, g[:, 2] / np.where(gmod < eps, 1.0, gmod))
    sent = np.sqrt(np.maximum(0.0, 1.0 - cost * cost))
    ylm[:, 0] = 1.0
    ylm[:, 1] = cost
    ylm[:, 3] = -sent / math.sqrt(2.0)
    for l in range(2, lmax + 1):
        for m in range(0, l - 1):
            lm, lm1, lm2 = l * l + 2 * m, (l - 1) ** 2 + 2 * m, (l - 2) ** 2 + 2 * m
            ylm[:, lm] = cost * (2 * l - 1) / math.sqrt(float(l * l - m * m)) * ylm[:, lm1] - math.sqrt(
                float((l - 1) * (l - 1) - m * m)) / math.sqrt(float(l * l - m * m)) * ylm[:, lm2]
        lm, lm1, lm2 = l * l + 2 * l, l * l + 2 * (l - 1), (l - 1) ** 2 + 2 * (l - 1)
        ylm[:, lm1] = cost * math.sqrt(float(2 * l - 1)) * ylm[:, lm2]
        ylm[:, lm] = -math.sqrt(float(2 * l - 1)) / math.sqrt(float(2 * l)) * sent * ylm[:, lm2]
    g1, g2 = g[:, 0], g[:, 1]
    with np.errstate(divide="ignore", invalid="ignore"):
        at = np.arctan(g2 / g1)
    phi = np.where(g1 > eps, at, np.where(g1 < -eps, at + math.pi, np.copysign(math.pi / 2.0, g2)))
    ylm[:, 0] = ylm[:, 0] / math.sqrt(FPI)
    lm = 0
    for l in range(1, lmax + 1):
        c = math.sqrt(float(2 * l + 1) / FPI)
        lm += 1
        ylm[:, lm] = c * ylm[:, lm]
        for m in range(1, l + 1):
            lm += 2
            q = ylm[:, lm].copy()
            ylm[:, lm - 1] = c * math.sqrt(2.0) * q * np.cos(m * phi)
            ylm[:, lm] = c * math.sqrt(2.0) * q * np.sin(m * phi)
    return ylm


def interp_beta(tab_beta_nt, dq, qg):
    """tab_beta_nt (nbetam, nqx) = Fortran tab_beta(:, :, nt) memory; qg (npw,) -> vq (npw, nbetam).
    beta_mod.f90:147-170."""
    nbetam, nqx = tab_beta_nt.shape
    t = qg / dq
    it = t.astype(np.int64)
    px = t - it
    ux, vx, wx = 1.0 - px, 2.0 - px, 3.0 - px
    inside = (it + 4) <= nqx
    i0 = np.where(inside, it, 0)
    vq = np.zeros((qg.size, nbetam))
    for nb in range(nbetam):
        tb = tab_beta_nt[nb]
        v = (tb[i0] * ux * vx * wx / 6.0 + tb[i0 + 1] * px * vx * wx / 2.0 - tb[i0 + 2] * px * ux * wx / 2.0
             + tb[i0 + 3] * px * ux * vx / 6.0)
        vq[:, nb] = np.where(inside, v, 0.0)
    return vq


def init_us_2(npw, npwx, igk, q, tau, ityp, tpiba, nr, eig, mill, g, us):
    """Returns vkb (nkb, npwx) (= Fortran vkb(npwx,nkb) memory). igk/ityp 1-based; eig = eigts(...);
    us: dict(nh (nsp,), nhtol/nhtolm/indv (nsp,nhm) [= Fortran (nhm,nsp) memory], lmaxkb, dq,
    tab_beta (nsp,nbetam,nqx)). init_us_2_acc.f90:73-160."""
    nh, lmaxkb = us["nh"], us["lmaxkb"]
    nsp = len(nh)
    nat = len(ityp)
    nkb = int(sum(nh[t - 1] for t in ityp))
    vkb = np.zeros((nkb, npwx), dtype=np.complex128)
    if lmaxkb < 0:
        return vkb
    iv = np.asarray(igk[:npw]) - 1
    gk = g[iv] + np.asarray(q)[None, :]
    qg2 = gk[:, 0] * gk[:, 0] + gk[:, 1] * gk[:, 1] + gk[:, 2] * gk[:, 2]
    ylm = ylmr2((lmaxkb + 1) ** 2, gk, qg2)
    qg = np.sqrt(qg2) * tpiba
    jkb = 0
    for nt in range(1, nsp + 1):
        vq = interp_beta(us["tab_beta"][nt - 1], us["dq"], qg)
        nhnt = int(nh[nt - 1])
        vkb1 = np.zeros((nhnt, npw))
        for ih in range(nhnt):
            nb = us["indv"][nt - 1, ih] - 1
            lm = us["nhtolm"][nt - 1, ih] - 1
            vkb1[ih] = ylm[:, lm] * vq[:, nb]
        for na in range(nat):
            if ityp[na] != nt:
                continue
            arg = (q[0] * tau[na, 0] + q[1] * tau[na, 1] + q[2] * tau[na, 2]) * TPI
            m = mill[iv]
            sk = (eig[0][na, m[:, 0] + nr[0]] * eig[1][na, m[:, 1] + nr[1]] * eig[2][na, m[:, 2] + nr[2]]
                  * complex(math.cos(arg), -math.sin(arg)))
            for ih in range(nhnt):
                pref = (-1j) ** int(us["nhtol"][nt - 1, ih])
                vkb[jkb + ih, :npw] = vkb1[ih] * sk * pref
            jkb += nhnt
    return vkb


def synthetic_uspp(nsp, lmaxkb, nbeta_per_l=2, nqx=400, dq=0.01, seed=0):
    """Random smooth tab_beta and the nh/nhtol/nhtolm/indv bookkeeping of an ultrasoft-like PP with nbeta_per_l radial
    betas for every l <= lmaxkb (the layout init_us_1 produces: for each beta nb of angular momentum l, 2l+1
    projectors with lm = l^2 + 1 .. (l+1)^2; upflib/init_us_1.f90 / pre_init)."""
    rng = np.random.default_rng(7234 + seed)
    nbetam = nbeta_per_l * (lmaxkb + 1)
    ls = [l for l in range(lmaxkb + 1) for _ in range(nbeta_per_l)]
    nh = np.full(nsp, sum(2 * l + 1 for l in ls), dtype=np.int32)
    nhm = int(nh.max())
    nhtol = np.zeros((nsp, nhm), np.int32)
    nhtolm = np.zeros((nsp, nhm), np.int32)
    indv = np.zeros((nsp, nhm), np.int32)
    for t in range(nsp):
        ih = 0
        for nb, l in enumerate(ls):
            for m in range(2 * l + 1):
                nhtol[t, ih], nhtolm[t, ih], indv[t, ih] = l, l * l + m + 1, nb + 1
                ih += 1
    x = np.arange(nqx) * dq
    tab = np.zeros((nsp, nbetam, nqx))
    for t in range(nsp):
        for nb, l in enumerate(ls):
            a, b, c = rng.uniform(0.5, 2.0), rng.uniform(0.3, 1.5), rng.uniform(0.0, 2.0)
            tab[t, nb] = a * x**l * np.exp(-b * x * x) * np.cos(c * x)
    return dict(nh=nh, nhm=nhm, nbetam=nbetam, lmaxkb=lmaxkb, nhtol=nhtol, nhtolm=nhtolm, indv=indv, nqx=nqx, dq=dq,
                tab_beta=tab)
