// Fused plane pass, second form ("plane2"): the transform of vloc_plane.cuh
//
//   G1[box][z][stick] --> plane(z) in shared memory: y DIF(+) | x DIF(+) * V(r)/N * x DIT(-) | y DIT(-) --> G1
//
// (the cft_2xy pair around the V(r) multiply, FFTXlib/src/fft_scalar.FFTW3.f90:197-365, PW/src/vloc_psi_acc.f90:197-229)
// with the same plane layout P[ypos*LDP + x] (LDP odd) and the same "lane = line" mapping, restructured after the ncu
// source page of k_vloc_plane_f (profiles/r2b0_*): there 19 % of the warp samples sat in the per-phase prologue (runtime
// divisions that deal butterflies to warps, a switch over the phase number), 25 % at the CTA barrier, and 48 % of the
// executed instructions were integer address arithmetic. Here
//   * every phase is a compile-time instantiation (template parameter PH): radix, strides, pass kind and the row
//     stride LDP are constants, so element addresses are one base register plus immediates;
//   * a phase's butterflies are a flat task list t = j*NG + g (NG line groups of 32 lanes) dealt round-robin to the
//     warps; the (g, j) of a warp's first task and its increment are computed ONCE per kernel (P2Roles), not per phase;
//   * the zero columns of the first / last x pass are a warp-uniform bit mask per butterfly (xmask), not a table
//     look-up per element;
//   * the kernel (vloc_plane2_inst.cuh) is persistent over a flat list of (box, z) planes and moves the stick rows with
//     bulk asynchronous copies (cp.async.bulk + mbarrier).
// Executed on the CPU by tests/emul (plane2_phase dispatches on the runtime phase number there).
#pragma once
#include "vloc_plane.cuh"

namespace qeb {

// per-warp task walk of one direction: first task (g0, j0) and the increment (dg, dj) for a stride of nw warps
struct P2Walk {
    int ng;          // line groups
    int g0, j0;      // task w
    int dg, dj;      // + nw tasks
    QEB_HD static P2Walk make(int w, int nw, int ng)
    {
        P2Walk k;
        k.ng = ng;
        k.g0 = w % ng;
        k.j0 = w / ng;
        k.dg = nw % ng;
        k.dj = nw / ng;
        return k;
    }
};
struct P2Roles {
    P2Walk y, x;
};

// bit b of xmask[a] (a < A0 = N/r0): column a + A0*b of the natural x axis is an active column
template <class PLX>
QEB_HD void plane2_xmask(const int* xmap, int* xmask)
{
    constexpr int R = PLX::r(0), A = PLX::N / R;
    for (int a = 0; a < A; ++a) {
        int m = 0;
        for (int b = 0; b < R; ++b)
            if (xmap[a + A * b] >= 0) m |= 1 << b;
        xmask[a] = m;
    }
}

// ---- one butterfly of a y pass for lane's column -------------------------------------------------------------------
// ymapT holds, per (y, xa), the stick id or - for a column position without a stick - the index of a spare element
// behind the row (stin: a zero; stout: a sink), so the pruned first / last pass needs no test per element.
template <class PL, int P, bool DIT, int LDP>
QEB_HD void p2_bfly_y(int j, bool active, int lane_col, int nxw, const cplx* tw, cplx* col, const cplx* stin, cplx* stout,
                      const int* ymapT)
{
    constexpr int R = PL::r(P), L = PL::L(P), A = L / R, NB = PL::N / L;
    constexpr double sgn = DIT ? -1.0 : +1.0;
    if (!active) return;
    const int blk = j / A, a = j - blk * A, base = blk * L + a;
    cplx v[R];
    if (!DIT && P == 0) {  // natural y in: the plane's sticks, zero elsewhere (pruned input)
        const int* ym = ymapT + (base * nxw + lane_col);
        const int ys = A * nxw;
#pragma unroll
        for (int b = 0; b < R; ++b) v[b] = stin[ym[b * ys]];
    } else {
        const cplx* c = col + base * LDP;
#pragma unroll
        for (int b = 0; b < R; ++b) v[b] = c[(A * b) * LDP];
    }
    if (DIT && A > 1 && a != 0) {  // a == 0: every twiddle is 1 (warp-uniform test)
        const int st = a * NB;
#pragma unroll
        for (int b = 1; b < R; ++b) v[b] = cmul(v[b], stw(tw, st * b, sgn));
    }
    Bfly<R>::run(v, sgn);
    if (!DIT && A > 1 && a != 0) {
        const int st = a * NB;
#pragma unroll
        for (int d = 1; d < R; ++d) v[d] = cmul(v[d], stw(tw, st * d, sgn));
    }
    if (DIT && P == 0) {  // natural y out: keep the stick rows
        const int* ym = ymapT + (base * nxw + lane_col);
        const int ys = A * nxw;
#pragma unroll
        for (int d = 0; d < R; ++d) stout[ym[d * ys]] = v[d];
    } else {
        cplx* c = col + base * LDP;
#pragma unroll
        for (int d = 0; d < R; ++d) c[(A * d) * LDP] = v[d];
    }
}

// ---- one butterfly of an x pass for lane's row ---------------------------------------------------------------------
// KIND 0: DIF pass P; 1: fused last DIF pass * V * first DIT pass (A = 1); 2: DIT pass P
template <class PL, int P, int KIND, int NY>
QEB_HD void p2_bfly_x(int j, bool active, const cplx* tw, cplx* row, const int* xmask, const double* vrow)
{
    constexpr int R = PL::r(P), L = PL::L(P), A = L / R, NB = PL::N / L;
    constexpr bool DIT = KIND == 2;
    constexpr double sgn = DIT ? -1.0 : +1.0;
    if (!active) return;
    const int blk = j / A, a = j - blk * A, base = blk * L + a;
    cplx v[R];
    double f[KIND == 1 ? R : 1];
    if (KIND == 1) {  // the pointwise factors first: their L2 latency hides behind the shared-memory loads + butterfly
#pragma unroll
        for (int d = 0; d < R; ++d) {
#if defined(__CUDA_ARCH__)
            f[d] = __ldg(vrow + (base + d) * NY);
#else
            f[d] = vrow[(base + d) * NY];
#endif
        }
    }
    cplx* r = row + base;
    if (KIND == 0 && P == 0 && PL::NP > 1) {  // natural x in: inactive columns are zero (warp-uniform mask)
        const int m = xmask[a];
#pragma unroll
        for (int b = 0; b < R; ++b) v[b] = ((m >> b) & 1) ? r[A * b] : mk(0.0, 0.0);
    } else if (KIND == 1 && PL::NP == 1) {  // single-pass plan: the fused pass reads natural x
        const int m = xmask[0];
#pragma unroll
        for (int b = 0; b < R; ++b) v[b] = ((m >> b) & 1) ? r[A * b] : mk(0.0, 0.0);
    } else {
#pragma unroll
        for (int b = 0; b < R; ++b) v[b] = r[A * b];
    }
    if (DIT && A > 1 && a != 0) {  // a == 0: every twiddle is 1 (warp-uniform test)
        const int st = a * NB;
#pragma unroll
        for (int b = 1; b < R; ++b) v[b] = cmul(v[b], stw(tw, st * b, sgn));
    }
    Bfly<R>::run(v, sgn);
    if (KIND == 1) {
#pragma unroll
        for (int d = 0; d < R; ++d) v[d] = scal(v[d], f[d]);
        Bfly<R>::run(v, -sgn);
    }
    if (KIND == 0 && A > 1 && a != 0) {
        const int st = a * NB;
#pragma unroll
        for (int d = 1; d < R; ++d) v[d] = cmul(v[d], stw(tw, st * d, sgn));
    }
    if ((DIT && P == 0) || (KIND == 1 && PL::NP == 1)) {  // natural x out: active columns only
        const int m = xmask[a];
#pragma unroll
        for (int d = 0; d < R; ++d)
            if ((m >> d) & 1) r[A * d] = v[d];
    } else {
#pragma unroll
        for (int d = 0; d < R; ++d) r[A * d] = v[d];
    }
}

template <class PLX, class PLY>
struct Plane2 {
    static constexpr int NX = PLX::N, NY = PLY::N;
    static constexpr int LDP = NX | 1;
    static constexpr int NPX = PLX::NP, NPY = PLY::NP;
    static constexpr int NPH = 2 * NPY + 2 * NPX - 1;
    static constexpr int NGX = (NY + 31) / 32;  // line groups of the x passes (rows)
    static constexpr int XMASK = PLX::N / PLX::r(0);
};

struct Plane2Sm {  // shared-memory tables of the plane2 kernel (host arrays in the CPU emulation)
    const cplx* twx;
    const cplx* twy;
    const int* ymapT;  // [nr2*nxw] stick id at (y, xa), or the spare element nst (plane2_ymap) where there is none
    const int* xact;   // [nxw]     xa -> x
    const int* xmask;  // [Plane2::XMASK]
};

// Phase PH of plane z for thread tid (lane = tid % 32 owns a line, warp = tid / 32 walks the task list).
// vz = pl_v + z*nr1*nr2 (the plane's pointwise factors, [xpos][ypos]).
template <class PLX, class PLY, int PH>
QEB_HD void plane2_phase_c(int tid, int nthr, int nxw, const P2Roles& ro, cplx* P, const cplx* stin, cplx* stout,
                           const Plane2Sm& sm, const double* vz)
{
    using G = Plane2<PLX, PLY>;
    constexpr int npy = G::NPY, npx = G::NPX, LDP = G::LDP;
    const int lane = tid & 31, nw = nthr >> 5;
    constexpr bool ypass = PH < npy || PH >= npy + 2 * npx - 1;
    if constexpr (ypass) {
        constexpr bool DIT = PH >= npy;
        constexpr int p = DIT ? (npy - 1 - (PH - (npy + 2 * npx - 1))) : PH;
        constexpr int NJ = PLY::N / PLY::r(p);
        const P2Walk& k = ro.y;
        const int ntask = NJ * k.ng;
        int g = k.g0, j = k.j0;
        for (int t = tid >> 5; t < ntask; t += nw) {
            const int xa = g * 32 + lane;
            const bool active = xa < nxw;
            cplx* col = P + sm.xact[active ? xa : 0];
            p2_bfly_y<PLY, p, DIT, LDP>(j, active, xa, nxw, sm.twy, col, stin, stout, sm.ymapT);
            g += k.dg;
            j += k.dj;
            if (g >= k.ng) {
                g -= k.ng;
                ++j;
            }
        }
    } else {
        constexpr int q = PH - npy;
        constexpr int KIND = q < npx - 1 ? 0 : (q == npx - 1 ? 1 : 2);
        constexpr int p = KIND == 0 ? q : (KIND == 1 ? npx - 1 : 2 * npx - 2 - q);
        constexpr int NJ = PLX::N / PLX::r(p);
        constexpr int NG = G::NGX;
        const P2Walk& k = ro.x;
        int g = k.g0, j = k.j0;
        for (int t = tid >> 5; t < NJ * NG; t += nw) {
            const int y = g * 32 + lane;
            const bool active = y < G::NY;
            cplx* row = P + (active ? y : 0) * LDP;
            p2_bfly_x<PLX, p, KIND, G::NY>(j, active, sm.twx, row, sm.xmask, vz + (active ? y : 0));
            g += k.dg;
            j += k.dj;
            if (g >= NG) {
                g -= NG;
                ++j;
            }
        }
    }
}

// runtime phase number -> compile-time instantiation (CPU emulation; the kernel unrolls the phase list itself)
template <class PLX, class PLY, int PH = 0>
struct Plane2Sel {
    QEB_HD static void run(int ph, int tid, int nthr, int nxw, const P2Roles& ro, cplx* P, const cplx* stin, cplx* stout,
                           const Plane2Sm& sm, const double* vz)
    {
        if (ph == PH)
            plane2_phase_c<PLX, PLY, PH>(tid, nthr, nxw, ro, P, stin, stout, sm, vz);
        else if constexpr (PH + 1 < Plane2<PLX, PLY>::NPH)
            Plane2Sel<PLX, PLY, PH + 1>::run(ph, tid, nthr, nxw, ro, P, stin, stout, sm, vz);
    }
};

QEB_HD P2Roles plane2_roles(int tid, int nthr, int nxw, int ngx)
{
    P2Roles r;
    const int w = tid >> 5, nw = nthr >> 5;
    r.y = P2Walk::make(w, nw, (nxw + 31) / 32);
    r.x = P2Walk::make(w, nw, ngx);
    return r;
}

// emulation entry: one phase for one thread, roles recomputed (the kernel keeps them in registers)
template <class PLX, class PLY>
QEB_HD void plane2_phase(int ph, int tid, int nthr, const VlocTables& t, int z, cplx* P, const cplx* stin, cplx* stout,
                         const Plane2Sm& sm)
{
    const P2Roles ro = plane2_roles(tid, nthr, t.nxw, Plane2<PLX, PLY>::NGX);
    Plane2Sel<PLX, PLY>::run(ph, tid, nthr, t.nxw, ro, P, stin, stout, sm, t.pl_v + (size_t)z * t.nr1 * t.nr2);
}

// stick rows carry spare elements behind the nst sticks (index nst: the zero of stin / the sink of stout)
QEB_HD int plane2_row_elems(int nst) { return (nst + 1 + 3) & ~3; }
QEB_HD int plane2_ymap(int s, int nst) { return s >= 0 ? s : nst; }

// shared-memory bytes: plane, two stick rows, twiddles, ymapT, xact, xmask, mbarrier
template <class PLX, class PLY>
QEB_HD size_t plane2_smem_bytes(int nst, int nxw)
{
    using G = Plane2<PLX, PLY>;
    const size_t nstp = (size_t)plane2_row_elems(nst);
    const size_t c16 = (size_t)G::NY * G::LDP + 2 * nstp + G::NX + G::NY;
    const size_t i4 = (size_t)G::NY * nxw + nxw + G::XMASK;
    return c16 * 16 + ((i4 * 4 + 15) & ~(size_t)15) + 32;
}

}  // namespace qeb
