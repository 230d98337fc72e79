// Fast (compile-time plan, warp-mapped) variants of the five V_loc passes of
// vloc_kernels.cuh. Same data flow, tables, intermediates and phase lists; the
// FFT length/radix list of the pass direction is a template parameter (SPlan)
// and tiles have at most 32 lines (one per lane). Selected by vloc.cu when the
// grid dimension has an instantiated static plan, otherwise the generic
// kernels run. `tw` is the twiddle table of the pass direction (the kernels
// stage it in shared memory). Executed on the CPU by tests/emul like the
// generic phases.
#pragma once
#include <stdlib.h>
#include "fft_static.cuh"
#include "vloc_kernels.cuh"

namespace qeb {

// software-pipelined loads in the global-source (first DIT) passes: measured SLOWER (zfwd 66 -> 81 us per 15 boxes at
// AUSURF112, Si512 25.8 -> 28.8 us/band: the extra live butterfly spills at the 128-register budget); off by default
#ifdef QEB_USE_GPF
constexpr bool QEB_GPF = true;
#else
constexpr bool QEB_GPF = false;
#endif

// small per-CTA tables the fast kernels stage in shared memory (in the CPU emulation they point at the
// global tables): twiddles of the pass direction, the x/y pruning map, the digit-reversal table, the
// stick list of the z group
struct FastTabs {
    const cplx* tw;
    const int* map;     // x pass: xmap[nr1]; y passes: ymap row of the plane [nr2]
    const int* rev;     // y passes: rev_y[nr2]; z passes: rev_z[nr3]
    const int* sticks;  // z passes: stick ids of the group [TZ]
};

struct LaneTile {  // S[pos*ld + lane]
    cplx* s;
    int ld;
    QEB_HD cplx load(int lane, int pos) const { return s[pos * ld + lane]; }
    QEB_HD void store(int lane, int pos, cplx v) const { s[pos * ld + lane] = v; }
};

// ---------------------------------------------------------------------------
// P1 fast: scatter + z inverse
// ---------------------------------------------------------------------------
template <class PL, int SUB = 1>
QEB_HD void zinv_fast_phase(int ph, int tid, int nthr, const VlocTables& t, const VlocIO& io, int grp, int box, cplx* smem,
                                const FastTabs& ft)
{
    if (ph < 2) {  // zero + scatter: identical to the generic kernel
        zinv_scatter_phase(ph, tid, nthr, t, io, grp, box, smem);
        return;
    }
    const int p = ph - 2;
    const int warp = tid / QEB_TL, lane = tid % QEB_TL, nwarps = nthr / QEB_TL;
    const int nl = t.grp_nl[grp];
    LaneTile s{smem, t.TZ};
    NoMid nm;
    if (p == PL::NP - 1 && t.peer_on) {  // fused exchange: straight into the plane owners' buffers
        G1PeerDst d{t.peer_base, t.peer_zown, t.peer_zi0, ft.rev, ft.sticks, t.peer_nb, box, t.nst};
        SPassSel<PL, false, false, 0, false, SUB>::run(p, warp, nwarps, lane, nl, ft.tw, +1.0, s, d, nm);
    } else if (p == PL::NP - 1) {
        G1Dst d{t.g1 + (size_t)box * t.g1_box, ft.rev, ft.sticks, t.g1_zs};
        SPassSel<PL, false, false, 0, false, SUB>::run(p, warp, nwarps, lane, nl, ft.tw, +1.0, s, d, nm);
    } else {
        SPassSel<PL, false, false, 0, false, SUB>::run(p, warp, nwarps, lane, nl, ft.tw, +1.0, s, s, nm);
    }
}

// ---------------------------------------------------------------------------
// P5 fast: z forward + gather
// ---------------------------------------------------------------------------
// STAGE: the G1 slab of the group is already in the tile (zfwd_stage: S[irev[z]*TZ + line], all of it requested at once
// with cp.async before phase 0), so the first DIT pass reads shared memory. Without it the first pass loads G1 itself:
// every warp then exposes one DRAM round trip per butterfly it owns (ncu source page, profiles/r2e_*: 18 % of the
// kernel's samples on the first use of those loads).
template <class PL, int SUB = 1, bool STAGE = false>
QEB_HD void zfwd_fast_phase(int ph, int tid, int nthr, const VlocTables& t, const VlocIO& io, int grp, int box, cplx* smem,
                                const FastTabs& ft)
{
    constexpr int np = PL::NP;
    if (ph >= np) {  // gather: identical to the generic kernel
        zfwd_gather_phase(tid, nthr, t, io, grp, box, smem);
        return;
    }
    const int p = np - 1 - ph;
    const int warp = tid / QEB_TL, lane = tid % QEB_TL, nwarps = nthr / QEB_TL;
    const int nl = t.grp_nl[grp];
    LaneTile s{smem, t.TZ};
    NoMid nm;
    if (ph == 0 && !STAGE) {
        G1Dst g{t.g1 + (size_t)box * t.g1_box, ft.rev, ft.sticks, t.g1_zs};
        SPassSel<PL, true, false, 0, QEB_GPF, SUB>::run(p, warp, nwarps, lane, nl, ft.tw, -1.0, g, s, nm);
    } else {
        SPassSel<PL, true, false, 0, false, SUB>::run(p, warp, nwarps, lane, nl, ft.tw, -1.0, s, s, nm);
    }
}

// the slab copy of the STAGE form: element (z, line) of the group -> S[irev[z]*TZ + line] (irev: inverse of ft.rev).
// On the device the copies are cp.async (the kernel waits for all of them once); the CPU emulation copies.
QEB_HD void zfwd_stage(int tid, int nthr, const VlocTables& t, int grp, int box, cplx* smem, const int* irev, const int* sticks)
{
    const int TZ = t.TZ, nl = t.grp_nl[grp];
    const cplx* g = t.g1 + (size_t)box * t.g1_box;
    const int tot = t.nr3 * TZ;
    for (int i = tid; i < tot; i += nthr) {
        const int z = i / TZ, l = i - z * TZ;
        if (l >= nl) continue;
        cplx* dst = smem + irev[z] * TZ + l;
        const cplx* src = g + (size_t)z * t.g1_zs + sticks[l];
#if defined(__CUDA_ARCH__)
        const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(src));
#else
        *dst = *src;
#endif
    }
}

// ---------------------------------------------------------------------------
// P2 / P4 fast: y passes
// ---------------------------------------------------------------------------
template <class PL>
QEB_HD void yinv_fast_phase(int ph, int tid, int nthr, const VlocTables& t, int ztile, int xa, int box, cplx* smem,
                                const FastTabs& ft)
{
    if (ph == 0) {
        yinv_load_phase(tid, nthr, t, ztile, xa, box, smem);  // G1 slab -> compact rows
        return;
    }
    constexpr int np = PL::NP;
    const int TY = t.TY, ldc = TY + 1;
    const int z0 = ztile * TY;
    const int nl = (t.nr3 - z0) < TY ? (t.nr3 - z0) : TY;
    cplx* C = smem;
    cplx* W = smem + (size_t)t.nsmax * ldc;
    const int p = ph - 1;
    const int warp = tid / QEB_TL, lane = tid % QEB_TL, nwarps = nthr / QEB_TL;
    LaneTile w{W, TY};
    NoMid nm;
    cplx* g2 = t.g2 + (size_t)box * t.g2_box + (size_t)xa * t.npos;
    G2YDst gd{g2, ft.rev, t.nr3, z0};
    YSrcCompact cs{C, ft.map, ldc};
    if (np == 1)
        SPassSel<PL, false, false>::run(p, warp, nwarps, lane, nl, ft.tw, +1.0, cs, gd, nm);
    else if (p == 0)
        SPassSel<PL, false, false>::run(p, warp, nwarps, lane, nl, ft.tw, +1.0, cs, w, nm);
    else if (p == np - 1)
        SPassSel<PL, false, false>::run(p, warp, nwarps, lane, nl, ft.tw, +1.0, w, gd, nm);
    else
        SPassSel<PL, false, false>::run(p, warp, nwarps, lane, nl, ft.tw, +1.0, w, w, nm);
}

template <class PL>
QEB_HD void yfwd_fast_phase(int ph, int tid, int nthr, const VlocTables& t, int ztile, int xa, int box, cplx* smem,
                                const FastTabs& ft)
{
    constexpr int np = PL::NP;
    if (ph >= np) {
        yfwd_store_phase(tid, nthr, t, ztile, xa, box, smem);  // compact rows -> G1 slab
        return;
    }
    const int TY = t.TY, ldc = TY + 1;
    const int z0 = ztile * TY;
    const int nl = (t.nr3 - z0) < TY ? (t.nr3 - z0) : TY;
    cplx* C = smem;
    cplx* W = smem + (size_t)t.nsmax * ldc;
    const int p = np - 1 - ph;
    const int warp = tid / QEB_TL, lane = tid % QEB_TL, nwarps = nthr / QEB_TL;
    LaneTile w{W, TY};
    NoMid nm;
    cplx* g2 = t.g2 + (size_t)box * t.g2_box + (size_t)xa * t.npos;
    G2YDst gs{g2, ft.rev, t.nr3, z0};
    YDstCompact cd{C, ft.map, ldc};
    if (np == 1)
        SPassSel<PL, true, false>::run(p, warp, nwarps, lane, nl, ft.tw, -1.0, gs, cd, nm);
    else if (ph == 0)
        SPassSel<PL, true, false, 0, QEB_GPF>::run(p, warp, nwarps, lane, nl, ft.tw, -1.0, gs, w, nm);
    else if (p == 0)
        SPassSel<PL, true, false>::run(p, warp, nwarps, lane, nl, ft.tw, -1.0, w, cd, nm);
    else
        SPassSel<PL, true, false>::run(p, warp, nwarps, lane, nl, ft.tw, -1.0, w, w, nm);
}

// ---------------------------------------------------------------------------
// P3 fast: fused x inverse * V * x forward
// ---------------------------------------------------------------------------
template <class PL>
QEB_HD void xfus_fast_phase(int ph, int tid, int nthr, const VlocTables& t, int tile, int box, cplx* smem,
                                const FastTabs& ft)
{
    constexpr int np = PL::NP;
    const int TX = t.TX;
    const int p0 = tile * TX;
    const int nl = (t.npos - p0) < TX ? (t.npos - p0) : TX;
    const int warp = tid / QEB_TL, lane = tid % QEB_TL, nwarps = nthr / QEB_TL;
    LaneTile w{smem, TX};
    G2XSrc g{t.g2 + (size_t)box * t.g2_box, ft.map, t.npos, p0};
    VMid vm{t.vp, t.npos, p0};
    NoMid nm;
    if (ph < np - 1) {
        if (ph == 0)
            SPassSel<PL, false, false>::run(ph, warp, nwarps, lane, nl, ft.tw, +1.0, g, w, nm);
        else
            SPassSel<PL, false, false>::run(ph, warp, nwarps, lane, nl, ft.tw, +1.0, w, w, nm);
    } else if (ph == np - 1) {
        if (np == 1)
            SPassSel<PL, false, true>::run(np - 1, warp, nwarps, lane, nl, ft.tw, +1.0, g, g, vm);
        else
            SPassSel<PL, false, true>::run(np - 1, warp, nwarps, lane, nl, ft.tw, +1.0, w, w, vm);
    } else {
        const int p = 2 * np - 2 - ph;
        if (p == 0)
            SPassSel<PL, true, false>::run(p, warp, nwarps, lane, nl, ft.tw, -1.0, w, g, nm);
        else
            SPassSel<PL, true, false>::run(p, warp, nwarps, lane, nl, ft.tw, -1.0, w, w, nm);
    }
}

// ---------------------------------------------------------------------------
// P3 fast for spinors (vloc_kernels.cuh::xspin_phase with a compile-time plan): the mix stays in registers between the
// last DIF and the first DIT butterflies of the two components (fft_static.cuh::spass_spin), so the phase list is the
// scalar one (2*NP-1) with two tiles.
// ---------------------------------------------------------------------------
struct SpinMix {
    const double* vnc;  // four arrays in the vp layout
    size_t vs;          // nr1*npos
    int npos, p0;
    QEB_HD void fetch(int line, int xpos, double v[4]) const
    {
        const size_t iv = (size_t)xpos * npos + p0 + line;
#if defined(__CUDA_ARCH__)
        v[0] = __ldg(vnc + iv), v[1] = __ldg(vnc + vs + iv), v[2] = __ldg(vnc + 2 * vs + iv), v[3] = __ldg(vnc + 3 * vs + iv);
#else
        v[0] = vnc[iv], v[1] = vnc[vs + iv], v[2] = vnc[2 * vs + iv], v[3] = vnc[3 * vs + iv];
#endif
    }
    QEB_HD void apply(const double v[4], cplx& u, cplx& d) const
    {
        const double a = v[0] + v[3], b = v[0] - v[3], v2 = v[1], v3 = v[2];
        const cplx nu = mk(u.x * a + (d.x * v2 + d.y * v3), u.y * a + (d.y * v2 - d.x * v3));
        const cplx nd = mk(d.x * b + (u.x * v2 - u.y * v3), d.y * b + (u.y * v2 + u.x * v3));
        u = nu;
        d = nd;
    }
};

template <class PL>
QEB_HD void xspin_fast_phase(int ph, int tid, int nthr, const VlocTables& t, int tile, int pair, cplx* smem,
                             const FastTabs& ft)
{
    constexpr int np = PL::NP;
    const int TX = t.TX;
    const int p0 = tile * TX;
    const int nl = (t.npos - p0) < TX ? (t.npos - p0) : TX;
    const int warp = tid / QEB_TL, lane = tid % QEB_TL, nwarps = nthr / QEB_TL;
    const size_t tile_el = (size_t)t.nr1 * TX;
    LaneTile wu{smem, TX}, wd{smem + tile_el, TX};
    G2XSrc gu{t.g2 + (size_t)(2 * pair) * t.g2_box, ft.map, t.npos, p0};
    G2XSrc gd{t.g2 + (size_t)(2 * pair + 1) * t.g2_box, ft.map, t.npos, p0};
    NoMid nm;
    if (ph == np - 1) {
        SpinMix mix{t.vnc, (size_t)t.nr1 * t.npos, t.npos, p0};
        if (np == 1)
            spass_spin<PL>(warp, nwarps, lane, nl, +1.0, gu, gd, mix);
        else
            spass_spin<PL>(warp, nwarps, lane, nl, +1.0, wu, wd, mix);
    } else if (ph < np - 1) {
        if (ph == 0) {
            SPassSel<PL, false, false>::run(ph, warp, nwarps, lane, nl, ft.tw, +1.0, gu, wu, nm);
            SPassSel<PL, false, false>::run(ph, warp, nwarps, lane, nl, ft.tw, +1.0, gd, wd, nm);
        } else {
            SPassSel<PL, false, false>::run(ph, warp, nwarps, lane, nl, ft.tw, +1.0, wu, wu, nm);
            SPassSel<PL, false, false>::run(ph, warp, nwarps, lane, nl, ft.tw, +1.0, wd, wd, nm);
        }
    } else {
        const int p = 2 * np - 2 - ph;
        if (p == 0) {
            SPassSel<PL, true, false>::run(p, warp, nwarps, lane, nl, ft.tw, -1.0, wu, gu, nm);
            SPassSel<PL, true, false>::run(p, warp, nwarps, lane, nl, ft.tw, -1.0, wd, gd, nm);
        } else {
            SPassSel<PL, true, false>::run(p, warp, nwarps, lane, nl, ft.tw, -1.0, wu, wu, nm);
            SPassSel<PL, true, false>::run(p, warp, nwarps, lane, nl, ft.tw, -1.0, wd, wd, nm);
        }
    }
}

// ---------------------------------------------------------------------------
// density accumulation fast: x inverse of one box of the tile + |psi(r)|^2 weights into the lane-mapped accumulator
// (vloc_kernels.cuh::xacc_phase with a compile-time plan)
// ---------------------------------------------------------------------------
struct RhoAccLane {  // racc[pos*ld + lane]
    double* racc;
    int ld;
    double w1, w2;
    QEB_HD void store(int lane, int pos, cplx v) const { racc[pos * ld + lane] += w1 * (v.x * v.x) + w2 * (v.y * v.y); }
};

template <class PL>
QEB_HD void xacc_fast_phase(int ph, int tid, int nthr, const VlocTables& t, const VlocIO& io, const double* w, int tile,
                            int box, cplx* smem, double* racc, const FastTabs& ft)
{
    constexpr int np = PL::NP;
    const int TX = t.TX;
    const int p0 = tile * TX;
    const int nl = (t.npos - p0) < TX ? (t.npos - p0) : TX;
    const int warp = tid / QEB_TL, lane = tid % QEB_TL, nwarps = nthr / QEB_TL;
    LaneTile wk{smem, TX};
    G2XSrc g{t.g2 + (size_t)box * t.g2_box, ft.map, t.npos, p0};
    NoMid nm;
    if (ph == np - 1) {
        RhoAccLane acc{racc, TX, 0.0, 0.0};
        xacc_weights(io, w, box, &acc.w1, &acc.w2);
        if (np == 1)
            SPassSel<PL, false, false>::run(ph, warp, nwarps, lane, nl, ft.tw, +1.0, g, acc, nm);
        else
            SPassSel<PL, false, false>::run(ph, warp, nwarps, lane, nl, ft.tw, +1.0, wk, acc, nm);
    } else if (ph == 0) {
        SPassSel<PL, false, false>::run(ph, warp, nwarps, lane, nl, ft.tw, +1.0, g, wk, nm);
    } else {
        SPassSel<PL, false, false>::run(ph, warp, nwarps, lane, nl, ft.tw, +1.0, wk, wk, nm);
    }
}

}  // namespace qeb
