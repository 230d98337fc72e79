"""ORACLE (test infrastructure, not product code).

CPU restatement of the reference's R&G ("slab") distribution of the smooth FFT grid over nproc ranks
and of the distributed V_loc / calbec data flow built on it. Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline leg may import this module.

Follows (paths relative to /root/reference):
  FFTXlib/src/stick_base.f90:214-262  sticks_map_set      G-vectors per (i1,i2) column within a cutoff
  FFTXlib/src/stick_base.f90:266-314  sticks_map_index    first-seen numbering of the columns
  FFTXlib/src/stick_base.f90:317-405  sticks_sort_new     new sticks appended by decreasing length (hpsort: ties by index)
  FFTXlib/src/stick_base.f90:409-545  sticks_dist_new     greedy owner = rank with fewest G-vectors, then fewest sticks
  FFTXlib/src/stick_base.f90:549-601  get_sticks
  FFTXlib/src/fft_types.f90:947-1066  fft_type_init       wave sticks (gkcut) first, then smooth sticks (gcut)
  FFTXlib/src/fft_types.f90:491-943   fft_type_set        nr3p/i0r3p, isind, iss, ismap, nsw, nsp (slab: nproc2 = 1)
  FFTXlib/src/fft_ggen.f90:60-96      fft_set_nl          nl = 1 + n3 + (isind - 1) * nr3x   (stick-major)
  Modules/recvec_subs.f90:126-245     ggen                local G list = global |G|^2 order restricted to own sticks
  FFTXlib/src/fft_scatter_2d.f90:38-266 fft_scatter       pack / alltoall / unpack via ismap (wave: isgn = +-2)
  FFTXlib/src/fft_parallel_2d.f90:36-204 tg_cft3s         cft_1z -> scatter -> cft_2xy and back
  Modules/becmod.f90:395-399,480      mp_sum(becp) after the local GEMM
Parity status: unpinned by reference fixtures (the reference has no golden data for the distribution);
checked here by construction properties (every stick owned once, +-G pairs co-located for Gamma, counts)
and by equality of the distributed data flow with the serial oracle (tests/test_dist_cpu.py).
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

from .descriptor import EPS8, Descriptor, hpsort_eps


def _column_counts(nr, bg, gcut, gamma_only):
    """sticks_map_set: st[i1,i2] = number of G in the column with |G|^2 <= gcut (Gamma: half space)."""
    ub = [(nr[0] - 1) // 2, (nr[1] - 1) // 2, (nr[2] - 1) // 2]
    i = np.arange(-ub[0], ub[0] + 1)
    j = np.arange(-ub[1], ub[1] + 1)
    k = np.arange(-ub[2], ub[2] + 1)
    I, J, K = np.meshgrid(i, j, k, indexing="ij")
    amod = ((I * bg[0, 0] + J * bg[0, 1] + K * bg[0, 2]) ** 2 + (I * bg[1, 0] + J * bg[1, 1] + K * bg[1, 2]) ** 2
            + (I * bg[2, 0] + J * bg[2, 1] + K * bg[2, 2]) ** 2)
    inside = amod <= gcut
    if gamma_only:
        inside &= I >= 0
        inside &= ~((I == 0) & (J < 0))
        inside &= ~((I == 0) & (J == 0) & (K < 0))
    return inside.sum(axis=2).astype(np.int64), ub  # indexed [i1+ub1, i2+ub2]


class SticksMap:
    """The persistent sticks_map of the reference (stick_base.f90:28-53): numbering, order and owners survive from
    one get_sticks call to the next."""

    def __init__(self, nr, bg, gamma_only, nproc):
        self.nr, self.bg, self.gamma, self.nproc = nr, bg, gamma_only, nproc
        self.ub = [(nr[0] - 1) // 2, (nr[1] - 1) // 2]
        self.index = {}   # (i1,i2) -> stick number (1-based)
        self.cols = []    # stick number-1 -> (i1,i2)
        self.idx = []     # processing order: stick numbers
        self.stown = {}   # (i1,i2) -> owner (1-based)

    def get_sticks(self, gcut):
        st, ub = _column_counts(self.nr, self.bg, gcut, self.gamma)
        # sticks_map_index: j2 outer, j1 inner, non-negative indices first then the negative ones
        n1, n2 = 2 * ub[0] + 1, 2 * ub[1] + 1
        for j2 in range(n2):
            i2 = j2 if j2 <= ub[1] else -ub[1] + (j2 - ub[1]) - 1
            for j1 in range(n1):
                i1 = j1 if j1 <= ub[0] else -ub[0] + (j1 - ub[0]) - 1
                if st[i1 + ub[0], i2 + ub[1]] > 0 and (i1, i2) not in self.index:
                    self.cols.append((i1, i2))
                    self.index[(i1, i2)] = len(self.cols)
        ngc = np.array([st[c[0] + ub[0], c[1] + ub[1]] for c in self.cols], dtype=np.int64)
        # sticks_sort_new: not-yet-ordered sticks with ngc > 0, by decreasing length, ties by stick number
        have = set(self.idx)
        new = [mc for mc in range(1, len(self.cols) + 1) if mc not in have and ngc[mc - 1] > 0]
        if self.nproc > 1:
            new.sort(key=lambda mc: (-ngc[mc - 1], mc))
        self.idx.extend(new)
        # sticks_dist_new (slab: nyfft = 1, iproc(1,j) = j)
        ncp = np.zeros(self.nproc, dtype=np.int64)
        ngp = np.zeros(self.nproc, dtype=np.int64)
        for mc in self.idx:
            i1, i2 = self.cols[mc - 1]
            if self.gamma and (i1 < 0 or (i1 == 0 and i2 < 0)):
                continue
            n = ngc[mc - 1]
            if n > 0 and self.stown.get((i1, i2), 0) == 0:
                jj = 0
                for j in range(self.nproc):
                    if ngp[j] < ngp[jj] or (ngp[j] == ngp[jj] and ncp[j] < ncp[jj]):
                        jj = j
                self.stown[(i1, i2)] = jj + 1
            if n > 0:
                o = self.stown[(i1, i2)] - 1
                ncp[o] += 1
                ngp[o] += n
        if self.gamma:
            for mc in self.idx:
                n = ngc[mc - 1]
                if n < 1:
                    continue
                i1, i2 = self.cols[mc - 1]
                jj = self.stown.get((i1, i2), 0)
                if i1 == 0 and i2 == 0:
                    if jj > 0:
                        ngp[jj - 1] += n - 1
                elif jj > 0:
                    self.stown[(-i1, -i2)] = jj
                    ncp[jj - 1] += 1
                    ngp[jj - 1] += n
        # owner map of the sticks relevant at this cutoff
        owner = {}
        for mc in self.idx:
            if ngc[mc - 1] > 0:
                i1, i2 = self.cols[mc - 1]
                owner[(i1, i2)] = self.stown[(i1, i2)]
                if self.gamma:
                    owner[(-i1, -i2)] = self.stown[(i1, i2)]
        return owner, ncp, ngp


@dataclass
class RankDescriptor:
    """Per-rank view: what dffts + gvect + wvfct hold on rank `mype` of an R&G-distributed run."""
    mype: int
    ngm: int                 # local smooth G-vectors
    ig_l2g: np.ndarray       # (ngm,) 0-based index into the global G list
    nl: np.ndarray           # (ngm,) 1-based, stick-major: 1 + n3 + (isind-1)*nr3
    nlm: np.ndarray | None
    npw: int
    igk: np.ndarray          # (npw,) 1-based into the local G list
    g2kin: np.ndarray
    pw_l2g: np.ndarray       # (npw,) 0-based row of the serial psi (position in the serial igk order)
    gstart: int


@dataclass
class DistDescriptor:
    nproc: int
    nr1: int
    nr2: int
    nr3: int
    gamma_only: bool
    nr3p: np.ndarray         # planes per rank
    i0r3p: np.ndarray        # first plane of each rank
    nsw: np.ndarray          # wave sticks per rank
    nsp: np.ndarray          # smooth sticks per rank
    iss: np.ndarray          # offset of each rank's sticks in ismap
    ismap: np.ndarray        # 1-based plane position (x fastest) of every stick, rank by rank, wave sticks first
    isind_owner: np.ndarray  # (nr1*nr2,) +owner for wave sticks, -owner for other smooth sticks, 0 none
    ranks: list = field(default_factory=list)


def distribute(d: Descriptor, nproc: int) -> DistDescriptor:
    """fft_type_init(dffts, 'wave', gkcut, dual=gcutms/gkcut) + ggen + fft_set_nl + gk_sort on every rank."""
    nr = (d.nr1, d.nr2, d.nr3)
    if d.nr3 < nproc:
        raise ValueError("there are processes with no planes (fft_types.f90:633-635)")
    smap = SticksMap(nr, d.bg, d.gamma_only, nproc)
    stw, ncpw, ngpw = smap.get_sticks(d.extra["gkcut"])
    st, ncp, ngp = smap.get_sticks(d.extra["gcutms"])
    # fft_type_set ------------------------------------------------------------------------------
    base, rem = divmod(d.nr3, nproc)
    nr3p = np.array([base + (1 if i < rem else 0) for i in range(nproc)], dtype=np.int32)
    i0r3p = np.concatenate([[0], np.cumsum(nr3p)[:-1]]).astype(np.int32)
    isind = np.zeros(d.nr1 * d.nr2, dtype=np.int32)
    for (i1, i2), own in st.items():
        m1 = i1 if i1 >= 0 else i1 + d.nr1
        m2 = i2 if i2 >= 0 else i2 + d.nr2
        isind[m1 + m2 * d.nr1] = own if stw.get((i1, i2), 0) > 0 else -own
    iss = np.concatenate([[0], np.cumsum(ncp)[:-1]]).astype(np.int32)
    ismap = np.zeros(int(ncp.sum()), dtype=np.int32)
    cnt = np.zeros(nproc, dtype=np.int64)
    local_index = np.zeros((nproc, d.nr1 * d.nr2), dtype=np.int32)  # isind as seen by each rank
    for sign in (1, -1):  # wave sticks first, then the remaining smooth sticks
        for pos in range(d.nr1 * d.nr2):
            ip = isind[pos] * sign
            if ip > 0:
                cnt[ip - 1] += 1
                ismap[cnt[ip - 1] - 1 + iss[ip - 1]] = pos + 1
                local_index[ip - 1, pos] = cnt[ip - 1]
        if sign == 1:
            nsw = cnt.copy()
            assert np.array_equal(nsw, ncpw), (nsw, ncpw)
    assert np.array_equal(cnt, ncp), (cnt, ncp)
    dd = DistDescriptor(nproc=nproc, nr1=d.nr1, nr2=d.nr2, nr3=d.nr3, gamma_only=d.gamma_only, nr3p=nr3p,
                        i0r3p=i0r3p, nsw=nsw.astype(np.int32), nsp=ncp.astype(np.int32), iss=iss, ismap=ismap,
                        isind_owner=isind)
    # ggen (global sort) + fft_set_nl + gk_sort, rank by rank ----------------------------------------
    n1 = np.where(d.mill[:, 0] < 0, d.mill[:, 0] + d.nr1, d.mill[:, 0]).astype(np.int64)
    n2 = np.where(d.mill[:, 1] < 0, d.mill[:, 1] + d.nr2, d.mill[:, 1]).astype(np.int64)
    n3 = np.where(d.mill[:, 2] < 0, d.mill[:, 2] + d.nr3, d.mill[:, 2]).astype(np.int64)
    pos = n1 + n2 * d.nr1
    owner_of_g = np.abs(isind[pos])
    assert np.all(owner_of_g > 0)
    serial_row = {int(ig): r for r, ig in enumerate(d.igk)}  # global G (1-based) -> row of the serial psi
    gcutw = d.extra["gcutw"]
    knorm2 = float(np.sum(d.xk ** 2))
    for me in range(nproc):
        mine = np.nonzero(owner_of_g == me + 1)[0]
        li = local_index[me]
        nl = (1 + n3[mine] + (li[pos[mine]].astype(np.int64) - 1) * d.nr3).astype(np.int32)
        nlm = None
        if d.gamma_only:
            m1 = np.where(-d.mill[mine, 0] < 0, -d.mill[mine, 0] + d.nr1, -d.mill[mine, 0]).astype(np.int64)
            m2 = np.where(-d.mill[mine, 1] < 0, -d.mill[mine, 1] + d.nr2, -d.mill[mine, 1]).astype(np.int64)
            m3 = np.where(-d.mill[mine, 2] < 0, -d.mill[mine, 2] + d.nr3, -d.mill[mine, 2]).astype(np.int64)
            lm = li[m1 + m2 * d.nr1].astype(np.int64)
            assert np.all(lm > 0), "the -G stick must live on the same rank"
            nlm = (1 + m3 + (lm - 1) * d.nr3).astype(np.int32)
        g = d.g[mine]
        kg = g + d.xk[None, :]
        q = np.sum(kg ** 2, axis=1)
        q = np.where(q <= EPS8, 0.0, q)
        sel = np.nonzero(q <= gcutw)[0]
        gk = q[sel]
        igk = (sel + 1).astype(np.int32)
        if knorm2 > EPS8 and igk.size:
            gk, igk = hpsort_eps(gk, igk)
        kgs = g[igk - 1] + d.xk[None, :]
        g2kin = (kgs[:, 0] * kgs[:, 0] + kgs[:, 1] * kgs[:, 1] + kgs[:, 2] * kgs[:, 2]) * d.tpiba2
        pw_l2g = np.array([serial_row[int(mine[i - 1]) + 1] for i in igk], dtype=np.int64)
        gstart = 2 if (mine.size and d.gg[mine[0]] <= EPS8) else 1
        dd.ranks.append(RankDescriptor(mype=me, ngm=mine.size, ig_l2g=mine, nl=nl, nlm=nlm, npw=igk.size, igk=igk,
                                       g2kin=g2kin, pw_l2g=pw_l2g, gstart=gstart))
    return dd


# ----------------------------------------------------------------------------
# distributed data flow (all ranks simulated in one process)
# ----------------------------------------------------------------------------
def _scatter_to_planes(dd: DistDescriptor, sticks):
    """fft_scatter isgn=+2 (fft_scatter_2d.f90:96-176): sticks[p] is (nsw[p], nr3) -> planes[q] is (nr3p[q], nr1*nr2),
    zero outside the wave sticks."""
    planes = []
    for q in range(dd.nproc):
        pl = np.zeros((dd.nr3p[q], dd.nr1 * dd.nr2), dtype=np.complex128)
        z0 = dd.i0r3p[q]
        for p in range(dd.nproc):
            mc = dd.ismap[dd.iss[p]:dd.iss[p] + dd.nsw[p]] - 1
            pl[:, mc] = sticks[p][:, z0:z0 + dd.nr3p[q]].T
        planes.append(pl)
    return planes


def _gather_from_planes(dd: DistDescriptor, planes):
    """fft_scatter isgn=-2 (fft_scatter_2d.f90:196-262)."""
    sticks = []
    for p in range(dd.nproc):
        mc = dd.ismap[dd.iss[p]:dd.iss[p] + dd.nsw[p]] - 1
        s = np.zeros((dd.nsw[p], dd.nr3), dtype=np.complex128)
        for q in range(dd.nproc):
            z0 = dd.i0r3p[q]
            s[:, z0:z0 + dd.nr3p[q]] = planes[q][:, mc].T
        sticks.append(s)
    return sticks


def dist_vloc_psi(dd: DistDescriptor, m: int, psi_loc, v_slab, hpsi_loc):
    """vloc_psi_k / vloc_psi_gamma over an R&G-distributed grid: psi_loc[r] is (m, lda_r) with the rank's own
    plane waves, v_slab[r] is (nr3p[r], nr2, nr1) = the rank's planes of V(r). tg_cft3s order
    (fft_parallel_2d.f90:36-204): cft_1z, fft_scatter, cft_2xy | cft_2xy, fft_scatter, cft_1z. hpsi_loc INOUT."""
    P = dd.nproc
    nr1, nr2, nr3 = dd.nr1, dd.nr2, dd.nr3
    step = 2 if dd.gamma_only else 1
    b = 0
    while b < m:
        pair = dd.gamma_only and b + 1 < m
        sticks = []
        for r in range(P):
            rk = dd.ranks[r]
            f = np.zeros(int(dd.nsw[r]) * nr3, dtype=np.complex128)
            n = rk.npw
            if not dd.gamma_only:
                f[rk.nl[rk.igk - 1] - 1] = psi_loc[r][b, :n]
            else:
                inl, inlm = rk.nl[:n] - 1, rk.nlm[:n] - 1
                c1 = psi_loc[r][b, :n]
                if pair:
                    c2 = psi_loc[r][b + 1, :n]
                    f[inl] = c1 + 1j * c2
                    f[inlm] = np.conj(c1 - 1j * c2)
                else:
                    f[inl] = c1
                    f[inlm] = np.conj(c1)
            s = f.reshape(int(dd.nsw[r]), nr3)
            sticks.append(np.fft.ifft(s, axis=1) * nr3)          # cft_1z, isign=+2: exp(+i), unscaled
        planes = _scatter_to_planes(dd, sticks)
        for q in range(P):
            pl = planes[q].reshape(int(dd.nr3p[q]), nr2, nr1)
            pl = np.fft.ifft2(pl, axes=(1, 2)) * (nr1 * nr2)      # cft_2xy
            pl = pl * v_slab[q]
            pl = np.fft.fft2(pl, axes=(1, 2)) / (nr1 * nr2)       # cft_2xy isign<0: scaled 1/(nx ny)
            planes[q] = pl.reshape(int(dd.nr3p[q]), nr1 * nr2)
        sticks = _gather_from_planes(dd, planes)
        for r in range(P):
            rk = dd.ranks[r]
            n = rk.npw
            f = (np.fft.fft(sticks[r], axis=1) / nr3).reshape(-1)  # cft_1z isign<0: scaled 1/nz
            if not dd.gamma_only:
                hpsi_loc[r][b, :n] += f[rk.nl[rk.igk - 1] - 1]
            else:
                inl, inlm = rk.nl[:n] - 1, rk.nlm[:n] - 1
                if pair:
                    fp = (f[inl] + f[inlm]) * 0.5
                    fm = (f[inl] - f[inlm]) * 0.5
                    hpsi_loc[r][b, :n] += fp.real + 1j * fm.imag
                    hpsi_loc[r][b + 1, :n] += fp.imag - 1j * fm.real
                else:
                    hpsi_loc[r][b, :n] += f[inl]
        b += step if pair or not dd.gamma_only else 1
    return hpsi_loc


def dist_h_psi(dd: DistDescriptor, m: int, psi_loc, v_slab, vkb_loc=None, deeq=None, nh_at=None, ofsbeta=None):
    """h_psi on every rank of the R&G group (h_psi.f90:138-147 kinetic, V_loc above, calbec with
    mp_sum(becp, intra_bgrp_comm) (becmod.f90:395-399 / :327-331), add_vuspsi on the rank's own rows,
    Gamma: Im hpsi(G=0)=0 on the rank that owns G=0). Returns (hpsi_loc list, becp)."""
    from . import hpsi_ref as R

    P = dd.nproc
    hpsi = []
    for r in range(P):
        rk = dd.ranks[r]
        lda = psi_loc[r].shape[1]
        h = np.zeros((m, lda), dtype=np.complex128)
        h[:, :rk.npw] = rk.g2kin[None, :] * psi_loc[r][:, :rk.npw]
        hpsi.append(h)
    dist_vloc_psi(dd, m, psi_loc, v_slab, hpsi)
    becp = None
    if vkb_loc is not None and vkb_loc[0].shape[0] > 0:
        for r in range(P):
            rk = dd.ranks[r]
            n = rk.npw
            if dd.gamma_only:
                part = R.calbec_gamma(n, vkb_loc[r], psi_loc[r], m, rk.gstart) if n else 0.0
            else:
                part = R.calbec_k(n, vkb_loc[r], psi_loc[r], m) if n else 0.0
            becp = part if becp is None else becp + part            # mp_sum
        ps = R.apply_deeq(becp, deeq, nh_at, ofsbeta)
        for r in range(P):
            n = dd.ranks[r].npw
            if n == 0:
                continue
            if dd.gamma_only:
                R.add_vuspsi_gamma(n, m, vkb_loc[r], ps, hpsi[r])
            else:
                R.add_vuspsi_k(n, m, vkb_loc[r], ps, hpsi[r])
    if dd.gamma_only:
        for r in range(P):
            if dd.ranks[r].gstart == 2 and dd.ranks[r].npw:
                hpsi[r][:, 0] = hpsi[r][:, 0].real
    return hpsi, becp


def split_rows(dd: DistDescriptor, full: np.ndarray, pad: int = 0, fill=0.0):
    """serial (m, lda) array -> per-rank (m, npw_r + pad) arrays holding the rank's own plane waves."""
    out = []
    for rk in dd.ranks:
        a = np.full((full.shape[0], rk.npw + pad), fill, dtype=full.dtype)
        a[:, :rk.npw] = full[:, rk.pw_l2g]
        out.append(a)
    return out


def v_slabs(dd: DistDescriptor, v: np.ndarray):
    """serial v(nnr) x-fastest -> per-rank (nr3p, nr2, nr1) slabs (the rank's z planes)."""
    box = v.reshape(dd.nr3, dd.nr2, dd.nr1)
    return [box[dd.i0r3p[r]:dd.i0r3p[r] + dd.nr3p[r]] for r in range(dd.nproc)]
