// The five pencil passes of the local-potential application
//   hpsi(G) (+)= FFT^-1-scaled[ V(r) * FFT[ psi(G) ] ]
// (reference: PW/src/vloc_psi_acc.f90:148-266 (k), :9-145 (Gamma);
//  FFTXlib/src/fft_helper_subroutines.f90:317-409,447-511,644-715,751-795;
//  FFTXlib/src/fft_scalar.FFTW3.f90:498-694 for the pass order and scaling).
//
// The full FFT box is never materialised. Intermediates:
//   G1[box][z][stick]        z-transformed sticks   (nst*nr3 per box)
//   G2[box][xa][y][z]        active x-planes        (nxw*nr2*nr3 per box)
// Passes (one kernel each, grid.y = box index within the batch):
//   P1 zinv : psi --scatter--> sticks, z DIF(+)            -> G1
//   P2 yinv : G1 plane slab, y DIF(+)                      -> G2
//   P3 xfus : G2 line tile, x DIF(+), *V(r)/N, x DIT(-)    -> G2 (in place)
//   P4 yfwd : G2, y DIT(-), keep stick rows                -> G1
//   P5 zfwd : G1, z DIT(-), gather (+ kinetic term)        -> hpsi
// Every kernel body is a list of barrier-separated phases, each a function of
// (tid, nthreads) so that tests/emul can execute the same code serially on a
// CPU. The product path only runs them as CUDA kernels (vloc.cu).
#pragma once
#include "fft_core.cuh"

namespace qeb {

struct VlocTables {
    int nr1, nr2, nr3;
    int nst, nxw, npos;  // sticks, active x-planes, nr2*nr3
    FftPlan px, py, pz;
    // planes
    const int* xmap;      // [nr1]  x -> xa or -1
    const int* plane_s0;  // [nxw]  first stick of the plane
    const int* plane_ns;  // [nxw]  sticks in the plane
    const int* ymap;      // [nxw*nr2] y -> local stick index in plane, or -1
    // z-pass stick groups
    int TZ, ngrp;
    const int* grp_stick;     // [ngrp*TZ] stick id (packed at the front), -1 pad
    const int* grp_nl;        // [ngrp] valid sticks in the group
    const FastDiv* grp_div;   // [ngrp] division by grp_nl
    const int* grp_coef_off;  // [ngrp+1]
    const int* coef_ig;       // [ncoef] 0-based row of psi/hpsi
    const int* coef_slot;     // [ncoef] z*TZ + l      (the +G point, nl)
    const int* coef_slot2;    // [ncoef] Gamma only: the -G point (nlm)
    // y / x tiles
    int TY, nty;  // z-lines per y-pass tile, tiles per plane
    FastDiv divTY, divTYlast;
    int TX, ntx;  // (y,z)-lines per x-pass tile
    FastDiv divTX, divTXlast;
    int nsmax;    // max sticks in a plane
    // potential, permuted + pre-scaled: vp[xpos*npos + (y*nr3+z)] = v(x=rev_x[xpos],y,z)/(nr1 nr2 nr3)
    const double* vp;
    // workspaces
    cplx* g1;
    cplx* g2;
    size_t g1_box, g2_box;  // elements per box
    // z passes only: element stride between consecutive z of one stick in G1 (nst when G1 is [box][z][stick]; the
    // slab-distributed path uses [z][box][stick] = nboxes*nst with g1_box = nst, so that the z range owned by a
    // peer rank is one contiguous block of the batch)
    size_t g1_zs;
    // fused plane pass (vloc_plane.cuh), when the (x,y) plane fits in shared memory; pl_v == nullptr: not used
    const int* pl_ymapT;   // [nr2*nxw] stick id at (y, xa) or -1
    const int* pl_xact;    // [nxw]     xa -> x
    const double* pl_v;    // [nr3][nr1 xpos][nr2 ypos] = v(rev_x[xpos], rev_y[ypos], z)/(nr1 nr2 nr3)
    int pl_ldp;            // row stride of the shared-memory plane (odd, >= nr1)
    // Slab-distributed path, fused exchange (distfft.h): peer_on != 0 makes the LAST z-inverse pass store each (z, stick)
    // element straight into the exchange buffer R of the rank that owns plane z - over NVLink when that is a peer -
    // in the layout the owner's y pass reads: R_q[block of the sender][z - i0(q)][box][stick of the sender]. 16 sticks of
    // a z-pass tile are 256 contiguous bytes there. peer_base[q] already points at this rank's block in q's slice.
    int peer_on;
    int peer_nb;                 // boxes of this batch
    cplx* peer_base[16];         // [nproc]
    const int* peer_zown;        // [nr3] owner of plane z
    const int* peer_zi0;         // [nproc] first plane of rank q
    // r_on != 0: the y passes read / write the exchange buffer R itself (t.g1 = this slot's R slice), element
    // (box, zl, sg) at r_nb*nr3*r_pref[sg] + (zl*r_nb + box)*r_nsw[sg] + r_sidx[sg] (fft_scatter_2d.f90:123,226 unpack / pack
    // through ismap, fused into the pass instead of a repack kernel)
    int r_on, r_nb;
    const int* r_pref;           // [nst] sticks on the ranks below the owner of sg
    const int* r_nsw;            // [nst] sticks of the owner of sg
    const int* r_sidx;           // [nst] position of sg in its owner's list
    // noncollinear 2x2 potential (vloc_psi_nc): four arrays in the vp layout, consecutive, nr1*npos doubles each:
    // v(:,1) .. v(:,4) of PW/src/vloc_psi_acc.f90:291, each /(nr1 nr2 nr3); nullptr: not set
    const double* vnc;
};

struct VlocIO {
    const cplx* psi;  // column b at psi + b*lda
    cplx* hpsi;
    const double* g2kin;  // NULL: accumulate into hpsi (vloc_psi semantics); else hpsi = g2kin*psi + vloc
    int lda, n;
    int band0;   // first band of this batch
    int nbands;  // bands in this call (m)
    int gamma;   // 1: two bands per box
};

QEB_HD FastDiv pick_div(int nl, int T, const FastDiv& full, const FastDiv& last) { return nl == T ? full : last; }

// ---------------------------------------------------------------------------
// P1: scatter + z inverse.  block = (group, box).  smem: nr3*TZ
// ---------------------------------------------------------------------------
struct G1Dst {  // last z-DIF pass: position -> z = rev[pos], line -> stick
    cplx* g1;
    const int* rev;
    const int* sticks;
    size_t zs;
    QEB_HD void store(int line, int pos, cplx v) const { g1[(size_t)rev[pos] * zs + sticks[line]] = v; }
    QEB_HD cplx load(int line, int pos) const { return g1[(size_t)rev[pos] * zs + sticks[line]]; }
};

struct G1PeerDst {  // last z-DIF pass of the fused exchange: (z, stick) -> R of the owner of plane z (VlocTables::peer_*)
    cplx* const* base;
    const int* zown;
    const int* zi0;
    const int* rev;
    const int* sticks;
    int nb, box, nst_me;
    QEB_HD void store(int line, int pos, cplx v) const
    {
        const int z = rev[pos], q = zown[z];
        base[q][((size_t)(z - zi0[q]) * nb + box) * nst_me + sticks[line]] = v;
    }
};

QEB_HD int zinv_nphase(const VlocTables& t) { return 2 + t.pz.npass; }

// phases 0 (zero the tile) and 1 (scatter psi into the sticks), shared with the fast kernels
QEB_HD void zinv_scatter_phase(int ph, int tid, int nthr, const VlocTables& t, const VlocIO& io, int grp, int box, cplx* smem)
{
    const int TZ = t.TZ;
    if (ph == 0) {
        const int tot = t.nr3 * TZ;
        for (int i = tid; i < tot; i += nthr) smem[i] = mk(0.0, 0.0);
        return;
    }
    if (ph == 1) {
        const int c0 = t.grp_coef_off[grp], c1 = t.grp_coef_off[grp + 1];
        if (!io.gamma) {
            const cplx* p = io.psi + (size_t)(io.band0 + box) * io.lda;
            // four coefficients per thread in flight: the index loads, then the dependent psi gathers, then the stores
            // (one global round trip per stage instead of one per coefficient)
            constexpr int U = 4;
            for (int c = c0 + tid; c < c1; c += U * nthr) {
                int ig[U], sl[U];
                cplx v[U];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int cc = c + u * nthr;
                    ig[u] = cc < c1 ? t.coef_ig[cc] : -1;
                    sl[u] = cc < c1 ? t.coef_slot[cc] : 0;
                }
#pragma unroll
                for (int u = 0; u < U; ++u)
                    if (ig[u] >= 0) v[u] = p[ig[u]];
#pragma unroll
                for (int u = 0; u < U; ++u)
                    if (ig[u] >= 0) smem[sl[u]] = v[u];
            }
        } else {
            const int b1 = io.band0 + 2 * box;
            const bool pair = (b1 + 1) < io.nbands;
            const cplx* p1 = io.psi + (size_t)b1 * io.lda;
            const cplx* p2 = p1 + io.lda;
            for (int c = c0 + tid; c < c1; c += nthr) {
                const int ig = t.coef_ig[c];
                const int sp = t.coef_slot[c], sm = t.coef_slot2[c];
                const cplx a = p1[ig];
                cplx vp, vm;
                if (pair) {
                    const cplx b = p2[ig];
                    vp = mk(a.x - b.y, a.y + b.x);   // c1 + i c2
                    vm = mk(a.x + b.y, b.x - a.y);   // conj(c1 - i c2)
                } else {
                    vp = a;
                    vm = mk(a.x, -a.y);
                }
                if (sp != sm) smem[sp] = vp;
                smem[sm] = vm;  // the reference writes nlm second: it wins at G=0
            }
        }
        return;
    }
}

QEB_HD void zinv_phase(int ph, int tid, int nthr, const VlocTables& t, const VlocIO& io, int grp, int box, cplx* smem)
{
    const int TZ = t.TZ;
    const int nl = t.grp_nl[grp];
    if (ph < 2) {
        zinv_scatter_phase(ph, tid, nthr, t, io, grp, box, smem);
        return;
    }
    const int p = ph - 2;
    const FastDiv dnl = t.grp_div[grp];
    const PassDesc pd = pass_of(t.pz, p);
    SmemTile s{smem, TZ};
    NoMid nm;
    if (p == t.pz.npass - 1 && t.peer_on) {
        G1PeerDst d{t.peer_base, t.peer_zown, t.peer_zi0, t.pz.rev, t.grp_stick + (size_t)grp * TZ, t.peer_nb, box, t.nst};
        fft_pass<false, false>(tid, nthr, nl, dnl, pd, t.pz.tw, +1.0, s, d, nm);
    } else if (p == t.pz.npass - 1) {
        G1Dst d{t.g1 + (size_t)box * t.g1_box, t.pz.rev, t.grp_stick + (size_t)grp * TZ, t.g1_zs};
        fft_pass<false, false>(tid, nthr, nl, dnl, pd, t.pz.tw, +1.0, s, d, nm);
    } else {
        fft_pass<false, false>(tid, nthr, nl, dnl, pd, t.pz.tw, +1.0, s, s, nm);
    }
}

// ---------------------------------------------------------------------------
// P5: z forward + gather.  block = (group, box).  smem: nr3*TZ
// ---------------------------------------------------------------------------
QEB_HD int zfwd_nphase(const VlocTables& t) { return 1 + t.pz.npass; }
QEB_HD void zfwd_gather_phase(int tid, int nthr, const VlocTables& t, const VlocIO& io, int grp, int box, cplx* smem);

QEB_HD void zfwd_phase(int ph, int tid, int nthr, const VlocTables& t, const VlocIO& io, int grp, int box, cplx* smem)
{
    const int TZ = t.TZ;
    const int nl = t.grp_nl[grp];
    const int np = t.pz.npass;
    if (ph < np) {
        const int p = np - 1 - ph;  // DIT: reverse order
        const FastDiv dnl = t.grp_div[grp];
        const PassDesc pd = pass_of(t.pz, p);
        SmemTile s{smem, TZ};
        NoMid nm;
        if (ph == 0) {
            G1Dst g{t.g1 + (size_t)box * t.g1_box, t.pz.rev, t.grp_stick + (size_t)grp * TZ, t.g1_zs};
            fft_pass<true, false>(tid, nthr, nl, dnl, pd, t.pz.tw, -1.0, g, s, nm);
        } else {
            fft_pass<true, false>(tid, nthr, nl, dnl, pd, t.pz.tw, -1.0, s, s, nm);
        }
        return;
    }
    zfwd_gather_phase(tid, nthr, t, io, grp, box, smem);
}

// gather (the 1/N scaling is folded into vp) + kinetic term; shared with the fast kernels
QEB_HD void zfwd_gather_phase(int tid, int nthr, const VlocTables& t, const VlocIO& io, int grp, int box, cplx* smem)
{
    const int c0 = t.grp_coef_off[grp], c1 = t.grp_coef_off[grp + 1];
    if (!io.gamma) {
        const size_t off = (size_t)(io.band0 + box) * io.lda;
        constexpr int U = 4;  // as in the scatter: index loads, dependent loads and stores in three stages
        for (int c = c0 + tid; c < c1; c += U * nthr) {
            int ig[U], sl[U];
            cplx a[U];
            double g2[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int cc = c + u * nthr;
                ig[u] = cc < c1 ? t.coef_ig[cc] : -1;
                sl[u] = cc < c1 ? t.coef_slot[cc] : 0;
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (ig[u] < 0) continue;
                if (io.g2kin) {
                    a[u] = io.psi[off + ig[u]];
                    g2[u] = io.g2kin[ig[u]];
                } else {
                    a[u] = io.hpsi[off + ig[u]];
                    g2[u] = 1.0;
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (ig[u] < 0) continue;
                const cplx v = smem[sl[u]];
                io.hpsi[off + ig[u]] = io.g2kin ? mk(g2[u] * a[u].x + v.x, g2[u] * a[u].y + v.y) : a[u] + v;
            }
        }
    } else {
        const int b1 = io.band0 + 2 * box;
        const bool pair = (b1 + 1) < io.nbands;
        const size_t o1 = (size_t)b1 * io.lda, o2 = o1 + io.lda;
        for (int c = c0 + tid; c < c1; c += nthr) {
            const int ig = t.coef_ig[c];
            const cplx vp = smem[t.coef_slot[c]], vm = smem[t.coef_slot2[c]];
            cplx r1, r2;
            if (pair) {
                // fp = (vp+vm)/2, fm = (vp-vm)/2; out1 = (Re fp, Im fm); out2 = (Im fp, -Re fm)
                const cplx fp = mk(0.5 * (vp.x + vm.x), 0.5 * (vp.y + vm.y));
                const cplx fm = mk(0.5 * (vp.x - vm.x), 0.5 * (vp.y - vm.y));
                r1 = mk(fp.x, fm.y);
                r2 = mk(fp.y, -fm.x);
            } else {
                r1 = vp;
                r2 = mk(0.0, 0.0);
            }
            if (io.g2kin) {
                const double g2 = io.g2kin[ig];
                const cplx p = io.psi[o1 + ig];
                io.hpsi[o1 + ig] = mk(g2 * p.x + r1.x, g2 * p.y + r1.y);
                if (pair) {
                    const cplx q = io.psi[o2 + ig];
                    io.hpsi[o2 + ig] = mk(g2 * q.x + r2.x, g2 * q.y + r2.y);
                }
            } else {
                io.hpsi[o1 + ig] = io.hpsi[o1 + ig] + r1;
                if (pair) io.hpsi[o2 + ig] = io.hpsi[o2 + ig] + r2;
            }
        }
    }
}

// ---------------------------------------------------------------------------
// P2: y inverse.  block = (z-tile, plane, box).
// smem: C[nsmax*(TY+1)] compact stick rows, then W[nr2*TY]
// ---------------------------------------------------------------------------
struct YSrcCompact {  // first y-DIF pass: y -> stick row of the plane (pruned)
    const cplx* c;
    const int* ymap;  // this plane's row
    int ld;
    QEB_HD cplx load(int line, int y) const
    {
        const int s = ymap[y];
        return s >= 0 ? c[s * ld + line] : mk(0.0, 0.0);
    }
};
struct G2YDst {  // y-pass <-> G2[xa][y = rev[pos]][z0+line]
    cplx* g2;    // at (box, xa)
    const int* rev;
    int nr3, z0;
    QEB_HD void store(int line, int pos, cplx v) const { g2[(size_t)rev[pos] * nr3 + z0 + line] = v; }
    QEB_HD cplx load(int line, int pos) const { return g2[(size_t)rev[pos] * nr3 + z0 + line]; }
};
struct YDstCompact {  // last y-DIT pass: natural y -> stick row (others dropped)
    cplx* c;
    const int* ymap;
    int ld;
    QEB_HD void store(int line, int y, cplx v) const
    {
        const int s = ymap[y];
        if (s >= 0) c[s * ld + line] = v;
    }
};

QEB_HD int ypass_nphase(const VlocTables& t) { return 1 + t.py.npass; }

// G1[box][z0+zl][s0+s] -> C[s][zl], lanes along s (shared with the fast kernels)
QEB_HD void yinv_load_phase(int tid, int nthr, const VlocTables& t, int ztile, int xa, int box, cplx* smem)
{
    const int TY = t.TY, ldc = TY + 1;
    const int z0 = ztile * TY;
    const int nl = (t.nr3 - z0) < TY ? (t.nr3 - z0) : TY;
    const int ns = t.plane_ns[xa], s0 = t.plane_s0[xa];
    cplx* C = smem;
    const cplx* g1 = t.g1 + (size_t)box * t.g1_box;
    const int lanes = nthr < 32 ? nthr : 32;
    const int nw = nthr / lanes, wid = tid / lanes, ln = tid % lanes;
    if (t.r_on) {  // straight from the exchange buffer (no unpack pass)
        for (int zl = wid; zl < nl; zl += nw)
            for (int s = ln; s < ns; s += lanes) {
                const int sg = s0 + s;
                C[s * ldc + zl] = t.g1[(size_t)t.r_nb * t.nr3 * t.r_pref[sg] + ((size_t)(z0 + zl) * t.r_nb + box) * t.r_nsw[sg] +
                                       t.r_sidx[sg]];
            }
        return;
    }
    for (int zl = wid; zl < nl; zl += nw)
        for (int s = ln; s < ns; s += lanes) C[s * ldc + zl] = g1[(size_t)(z0 + zl) * t.nst + s0 + s];
}

// C[s][zl] -> G1[box][z0+zl][s0+s] (shared with the fast kernels)
QEB_HD void yfwd_store_phase(int tid, int nthr, const VlocTables& t, int ztile, int xa, int box, cplx* smem)
{
    const int TY = t.TY, ldc = TY + 1;
    const int z0 = ztile * TY;
    const int nl = (t.nr3 - z0) < TY ? (t.nr3 - z0) : TY;
    const int ns = t.plane_ns[xa], s0 = t.plane_s0[xa];
    const cplx* C = smem;
    cplx* g1 = t.g1 + (size_t)box * t.g1_box;
    const int lanes = nthr < 32 ? nthr : 32;
    const int nw = nthr / lanes, wid = tid / lanes, ln = tid % lanes;
    if (t.r_on) {  // straight into the exchange buffer (no pack pass)
        for (int zl = wid; zl < nl; zl += nw)
            for (int s = ln; s < ns; s += lanes) {
                const int sg = s0 + s;
                t.g1[(size_t)t.r_nb * t.nr3 * t.r_pref[sg] + ((size_t)(z0 + zl) * t.r_nb + box) * t.r_nsw[sg] + t.r_sidx[sg]] =
                    C[s * ldc + zl];
            }
        return;
    }
    for (int zl = wid; zl < nl; zl += nw)
        for (int s = ln; s < ns; s += lanes) g1[(size_t)(z0 + zl) * t.nst + s0 + s] = C[s * ldc + zl];
}

QEB_HD void yinv_phase(int ph, int tid, int nthr, const VlocTables& t, int ztile, int xa, int box, cplx* smem)
{
    const int TY = t.TY, ldc = TY + 1;
    const int z0 = ztile * TY;
    const int nl = (t.nr3 - z0) < TY ? (t.nr3 - z0) : TY;
    cplx* C = smem;
    cplx* W = smem + (size_t)t.nsmax * ldc;
    const int np = t.py.npass;
    if (ph == 0) {
        yinv_load_phase(tid, nthr, t, ztile, xa, box, smem);
        return;
    }
    const int p = ph - 1;
    const FastDiv dnl = pick_div(nl, TY, t.divTY, t.divTYlast);
    const PassDesc pd = pass_of(t.py, p);
    SmemTile w{W, TY};
    NoMid nm;
    cplx* g2 = t.g2 + (size_t)box * t.g2_box + (size_t)xa * t.npos;
    G2YDst gd{g2, t.py.rev, t.nr3, z0};
    YSrcCompact cs{C, t.ymap + (size_t)xa * t.nr2, ldc};
    if (np == 1) {
        fft_pass<false, false>(tid, nthr, nl, dnl, pd, t.py.tw, +1.0, cs, gd, nm);
    } else if (p == 0) {
        fft_pass<false, false>(tid, nthr, nl, dnl, pd, t.py.tw, +1.0, cs, w, nm);
    } else if (p == np - 1) {
        fft_pass<false, false>(tid, nthr, nl, dnl, pd, t.py.tw, +1.0, w, gd, nm);
    } else {
        fft_pass<false, false>(tid, nthr, nl, dnl, pd, t.py.tw, +1.0, w, w, nm);
    }
}

// ---------------------------------------------------------------------------
// P4: y forward.  block = (z-tile, plane, box)
// ---------------------------------------------------------------------------
QEB_HD void yfwd_phase(int ph, int tid, int nthr, const VlocTables& t, int ztile, int xa, int box, cplx* smem)
{
    const int TY = t.TY, ldc = TY + 1;
    const int z0 = ztile * TY;
    const int nl = (t.nr3 - z0) < TY ? (t.nr3 - z0) : TY;
    cplx* C = smem;
    cplx* W = smem + (size_t)t.nsmax * ldc;
    const int np = t.py.npass;
    if (ph < np) {
        const int p = np - 1 - ph;
        const FastDiv dnl = pick_div(nl, TY, t.divTY, t.divTYlast);
        const PassDesc pd = pass_of(t.py, p);
        SmemTile w{W, TY};
        NoMid nm;
        cplx* g2 = t.g2 + (size_t)box * t.g2_box + (size_t)xa * t.npos;
        G2YDst gs{g2, t.py.rev, t.nr3, z0};
        YDstCompact cd{C, t.ymap + (size_t)xa * t.nr2, ldc};
        if (np == 1) {
            fft_pass<true, false>(tid, nthr, nl, dnl, pd, t.py.tw, -1.0, gs, cd, nm);
        } else if (ph == 0) {
            fft_pass<true, false>(tid, nthr, nl, dnl, pd, t.py.tw, -1.0, gs, w, nm);
        } else if (p == 0) {
            fft_pass<true, false>(tid, nthr, nl, dnl, pd, t.py.tw, -1.0, w, cd, nm);
        } else {
            fft_pass<true, false>(tid, nthr, nl, dnl, pd, t.py.tw, -1.0, w, w, nm);
        }
        return;
    }
    yfwd_store_phase(tid, nthr, t, ztile, xa, box, smem);
}

// ---------------------------------------------------------------------------
// P3: fused x inverse * V * x forward.  block = (line tile, box). smem: nr1*TX
// ---------------------------------------------------------------------------
struct G2XSrc {  // natural x -> active plane (pruned); lines = (y,z) positions
    cplx* g2;    // at box
    const int* xmap;
    int npos, p0;
    QEB_HD cplx load(int line, int x) const
    {
        const int xa = xmap[x];
        return xa >= 0 ? g2[(size_t)xa * npos + p0 + line] : mk(0.0, 0.0);
    }
    QEB_HD void store(int line, int x, cplx v) const
    {
        const int xa = xmap[x];
        if (xa >= 0) g2[(size_t)xa * npos + p0 + line] = v;
    }
};
struct VMid {
    const double* vp;
    int npos, p0;
    QEB_HD double fetch(int line, int xpos) const
    {
#if defined(__CUDA_ARCH__)
        return __ldg(vp + (size_t)xpos * npos + p0 + line);
#else
        return vp[(size_t)xpos * npos + p0 + line];
#endif
    }
    QEB_HD cplx apply(int line, int xpos, cplx v) const
    {
        const double s = fetch(line, xpos);
        return mk(v.x * s, v.y * s);
    }
};

QEB_HD int xfus_nphase(const VlocTables& t) { return 2 * t.px.npass - 1; }

QEB_HD void xfus_phase(int ph, int tid, int nthr, const VlocTables& t, int tile, int box, cplx* smem)
{
    const int TX = t.TX;
    const int p0 = tile * TX;
    const int nl = (t.npos - p0) < TX ? (t.npos - p0) : TX;
    const FastDiv dnl = pick_div(nl, TX, t.divTX, t.divTXlast);
    const int np = t.px.npass;
    SmemTile w{smem, TX};
    G2XSrc g{t.g2 + (size_t)box * t.g2_box, t.xmap, t.npos, p0};
    VMid vm{t.vp, t.npos, p0};
    NoMid nm;
    if (ph < np - 1) {  // DIF passes 0 .. np-2
        const PassDesc pd = pass_of(t.px, ph);
        if (ph == 0)
            fft_pass<false, false>(tid, nthr, nl, dnl, pd, t.px.tw, +1.0, g, w, nm);
        else
            fft_pass<false, false>(tid, nthr, nl, dnl, pd, t.px.tw, +1.0, w, w, nm);
    } else if (ph == np - 1) {  // last DIF pass * V * first DIT pass
        const PassDesc pd = pass_of(t.px, np - 1);
        if (np == 1)
            fft_pass<false, true>(tid, nthr, nl, dnl, pd, t.px.tw, +1.0, g, g, vm);
        else
            fft_pass<false, true>(tid, nthr, nl, dnl, pd, t.px.tw, +1.0, w, w, vm);
    } else {  // DIT passes np-2 .. 0
        const int p = 2 * np - 2 - ph;
        const PassDesc pd = pass_of(t.px, p);
        if (p == 0)
            fft_pass<true, false>(tid, nthr, nl, dnl, pd, t.px.tw, -1.0, w, g, nm);
        else
            fft_pass<true, false>(tid, nthr, nl, dnl, pd, t.px.tw, -1.0, w, w, nm);
    }
}

// ---------------------------------------------------------------------------
// P3 for spinors (noncollinear with magnetisation, PW/src/vloc_psi_acc.f90:315-341): the two components of a band are
// boxes 2*pair and 2*pair+1 of the batch. block = (line tile, pair); both components are x-transformed into two
// shared-memory tiles, mixed point by point with the 2x2 potential
//   sup  = psi_up (v1+v4) + psi_dn (v2 - i v3),   sdwn = psi_dn (v1-v4) + psi_up (v2 + i v3)
// and transformed back. Phases: np DIF passes | mix | np DIT passes. smem: 2*nr1*TX complex.
// ---------------------------------------------------------------------------
QEB_HD int xspin_nphase(const VlocTables& t) { return 2 * t.px.npass + 1; }

QEB_HD void xspin_phase(int ph, int tid, int nthr, const VlocTables& t, int tile, int pair, cplx* smem)
{
    const int TX = t.TX;
    const int p0 = tile * TX;
    const int nl = (t.npos - p0) < TX ? (t.npos - p0) : TX;
    const FastDiv dnl = pick_div(nl, TX, t.divTX, t.divTXlast);
    const int np = t.px.npass;
    const size_t tile_el = (size_t)t.nr1 * TX;
    NoMid nm;
    if (ph == np) {  // 2x2 mix in position order (the four potentials are permuted like vp)
        const size_t vs = (size_t)t.nr1 * t.npos;
        for (int i = tid; i < t.nr1 * nl; i += nthr) {
            const int xpos = i / nl, line = i - xpos * nl;
            const size_t iv = (size_t)xpos * t.npos + p0 + line;
            const double v1 = t.vnc[iv], v2 = t.vnc[vs + iv], v3 = t.vnc[2 * vs + iv], v4 = t.vnc[3 * vs + iv];
            cplx* pu = smem + xpos * TX + line;
            cplx* pd = pu + tile_el;
            const cplx u = *pu, d = *pd;
            const double a = v1 + v4, b = v1 - v4;
            // d*(v2 - i v3) = (d.x v2 + d.y v3) + i (d.y v2 - d.x v3);  u*(v2 + i v3) = (u.x v2 - u.y v3) + i (u.y v2 + u.x v3)
            *pu = mk(u.x * a + (d.x * v2 + d.y * v3), u.y * a + (d.y * v2 - d.x * v3));
            *pd = mk(d.x * b + (u.x * v2 - u.y * v3), d.y * b + (u.y * v2 + u.x * v3));
        }
        return;
    }
    for (int comp = 0; comp < 2; ++comp) {
        SmemTile w{smem + comp * tile_el, TX};
        G2XSrc g{t.g2 + (size_t)(2 * pair + comp) * t.g2_box, t.xmap, t.npos, p0};
        if (ph < np) {  // DIF pass ph
            const PassDesc pd = pass_of(t.px, ph);
            if (ph == 0)
                fft_pass<false, false>(tid, nthr, nl, dnl, pd, t.px.tw, +1.0, g, w, nm);
            else
                fft_pass<false, false>(tid, nthr, nl, dnl, pd, t.px.tw, +1.0, w, w, nm);
        } else {  // DIT pass np-1 .. 0
            const int p = 2 * np - ph;
            const PassDesc pd = pass_of(t.px, p);
            if (p == 0)
                fft_pass<true, false>(tid, nthr, nl, dnl, pd, t.px.tw, -1.0, w, g, nm);
            else
                fft_pass<true, false>(tid, nthr, nl, dnl, pd, t.px.tw, -1.0, w, w, nm);
        }
    }
}

// ---------------------------------------------------------------------------
// Density accumulation (sum_band, PW/src/sum_band.f90:376-394 Gamma, :578-593 k; get_rho_k :650-671,
// get_rho_gamma :674-696): P1 and P2 as above, then per line tile the x inverse transform of every box of the batch
// and rho(r) += w1 |psi(r)|^2 (k) / w1 Re^2 + w2 Im^2 (Gamma, two bands per box), accumulated in shared memory over
// the boxes and added to rho once per batch. block = line tile. smem: tile nr1*TX complex, racc nr1*TX doubles.
// ---------------------------------------------------------------------------
struct RhoAccDst {  // last x-DIF pass: position -> accumulator (each (pos,line) is owned by one thread)
    double* racc;
    int ld;
    double w1, w2;
    QEB_HD void store(int line, int pos, cplx v) const { racc[pos * ld + line] += w1 * (v.x * v.x) + w2 * (v.y * v.y); }
};

QEB_HD void xacc_weights(const VlocIO& io, const double* w, int box, double* w1, double* w2)
{
    if (!io.gamma) {
        *w1 = *w2 = w[io.band0 + box];
    } else {
        const int b1 = io.band0 + 2 * box;
        *w1 = w[b1];
        *w2 = (b1 + 1) < io.nbands ? w[b1 + 1] : w[b1];  // sum_band.f90:385-391
    }
}

QEB_HD void xacc_zero_phase(int tid, int nthr, const VlocTables& t, double* racc)
{
    for (int i = tid; i < t.nr1 * t.TX; i += nthr) racc[i] = 0.0;
}

QEB_HD void xacc_phase(int ph, int tid, int nthr, const VlocTables& t, const VlocIO& io, const double* w, int tile,
                       int box, cplx* smem, double* racc)
{
    const int TX = t.TX;
    const int p0 = tile * TX;
    const int nl = (t.npos - p0) < TX ? (t.npos - p0) : TX;
    const FastDiv dnl = pick_div(nl, TX, t.divTX, t.divTXlast);
    const int np = t.px.npass;
    SmemTile wk{smem, TX};
    G2XSrc g{t.g2 + (size_t)box * t.g2_box, t.xmap, t.npos, p0};
    NoMid nm;
    const PassDesc pd = pass_of(t.px, ph);
    if (ph == np - 1) {
        RhoAccDst acc{racc, TX, 0.0, 0.0};
        xacc_weights(io, w, box, &acc.w1, &acc.w2);
        if (np == 1)
            fft_pass<false, false>(tid, nthr, nl, dnl, pd, t.px.tw, +1.0, g, acc, nm);
        else
            fft_pass<false, false>(tid, nthr, nl, dnl, pd, t.px.tw, +1.0, wk, acc, nm);
    } else if (ph == 0) {
        fft_pass<false, false>(tid, nthr, nl, dnl, pd, t.px.tw, +1.0, g, wk, nm);
    } else {
        fft_pass<false, false>(tid, nthr, nl, dnl, pd, t.px.tw, +1.0, wk, wk, nm);
    }
}

// rho(x, y, z) += racc[pos(x)][line], line p0+l = y*nr3 + z; irev[x] = position holding x after the DIF
QEB_HD void xacc_flush_phase(int tid, int nthr, const VlocTables& t, int tile, const int* irev, const double* racc,
                             double* rho, bool shared_tile = false)
{
    const int TX = t.TX;
    const int p0 = tile * TX;
    const int nl = (t.npos - p0) < TX ? (t.npos - p0) : TX;
    for (int i = tid; i < nl * t.nr1; i += nthr) {
        const int line = i / t.nr1, x = i - line * t.nr1;
        const int p = p0 + line;
        const int y = p / t.nr3, z = p - y * t.nr3;
        double* r = rho + x + (size_t)t.nr1 * (y + (size_t)t.nr2 * z);
        const double v = racc[irev[x] * TX + line];
#if defined(__CUDA_ARCH__)
        if (shared_tile)  // several CTAs (box chunks) own this tile: order of the additions is not fixed
            atomicAdd(r, v);
        else
#endif
            *r += v;
    }
}

}  // namespace qeb
