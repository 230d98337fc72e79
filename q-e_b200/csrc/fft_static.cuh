// Compile-time FFT plans and the warp-mapped radix pass used by the fast V_loc
// kernels (vloc_fast.cuh). Same algorithm as fft_core.cuh (in-place DIF/DIT
// passes over positions base + A*b), but
//   * the length and the radix list are template parameters, so every stride,
//     sub-length and division is a constant and the pass list is unrolled;
//   * the work mapping is "one line per lane, one butterfly index per warp":
//     lane l owns line l of the tile (tiles have <= 32 lines), the warps stride
//     over the N/R butterfly indices j, so (blk, a), the twiddle addresses and
//     all table look-ups are warp-uniform and the shared-memory accesses
//     S[pos*32 + lane] are conflict-free.
#pragma once
#include "fft_core.cuh"

namespace qeb {

template <int N_, int... Rs>
struct SPlan {
    static constexpr int N = N_;
    static constexpr int NP = (int)sizeof...(Rs);
    QEB_HD static constexpr int r(int p)
    {
        constexpr int a[sizeof...(Rs)] = {Rs...};
        return a[p];
    }
    QEB_HD static constexpr int L(int p)
    {
        int l = N_;
        for (int i = 0; i < p; ++i) l /= r(i);
        return l;
    }
    // does a runtime plan (host factorisation) describe the same pass list?
    static bool matches(const FftPlan& pl)
    {
        if (pl.n != N_ || pl.npass != NP) return false;
        for (int p = 0; p < NP; ++p)
            if (pl.r[p] != r(p)) return false;
        return true;
    }
};

constexpr int QEB_TL = 32;  // lines per tile = lanes
#ifdef QEB_NO_VPREFETCH       // tuning aid (q-e_b200/build.py --variant): fetch the pointwise factor without pipelining
constexpr bool QEB_VPF = false;
#else
constexpr bool QEB_VPF = true;
#endif

// twiddle from a table that the fast kernels stage in shared memory (plain load: LDS once inlined)
QEB_HD cplx stw(const cplx* tw, int j, double sgn) { return mk(tw[j].x, sgn * tw[j].y); }

// One radix pass P of plan PL over the tile; see fft_core.cuh::fft_pass_r for the math.
// PF: software-pipelined input (the loads of the next butterfly are issued before the current one is computed) for
// passes whose source is global memory, where each warp otherwise exposes one L2/DRAM round trip per butterfly
// SUB: butterflies per warp. Tiles of at most 16 lines (the z passes at nr3 = 200: 51 KB per 16-stick tile) would leave half
// of every warp idle with one line per lane; with SUB = 2 lanes 0-15 and 16-31 carry the same 16 lines of two consecutive
// butterfly indices, so the FP64 instructions run with full warps and a warp keeps twice the loads in flight.
template <class PL, int P, bool DIT, bool FUSE2, bool PF = false, int SUB = 1, class Src, class Dst, class Mid>
QEB_HD void spass(int warp, int nwarps, int lane_, int nl, const cplx* tw, double sgn, Src& src, Dst& dst, Mid& mid)
{
    constexpr int R = PL::r(P);
    constexpr int L = PL::L(P);
    constexpr int A = L / R;
    constexpr int NB = PL::N / L;
    constexpr int NJ = PL::N / R;
    constexpr int TLN = QEB_TL / SUB;
    const int sub = SUB > 1 ? lane_ / TLN : 0;
    const int lane = SUB > 1 ? lane_ - sub * TLN : lane_;  // the line
    const int j0 = warp * SUB + sub, jstep = nwarps * SUB;
    if (lane >= nl) return;
    double fnext[FUSE2 ? R : 1];
    cplx vnext[PF ? R : 1];
    if (FUSE2 && QEB_VPF && j0 < NJ) {  // pointwise factor of the first butterfly (software pipeline, see below)
        const int blk = j0 / A;
        const int base = blk * L + (j0 - blk * A);
#pragma unroll
        for (int d = 0; d < R; ++d) fnext[d] = mid.fetch(lane, base + d);
    }
    for (int j = j0; j < NJ; j += jstep) {
        const int blk = j / A;
        const int a = j - blk * A;
        const int base = blk * L + a;
        cplx v[R];
        double f[FUSE2 ? R : 1];
        if (PF) {
            if (j == j0) {
#pragma unroll
                for (int b = 0; b < R; ++b) vnext[b] = src.load(lane, base + A * b);
            }
#pragma unroll
            for (int b = 0; b < R; ++b) v[b] = vnext[b];
            const int jn = j + jstep;
            if (jn < NJ) {
                const int blkn = jn / A;
                const int basen = blkn * L + (jn - blkn * A);
#pragma unroll
                for (int b = 0; b < R; ++b) vnext[b] = src.load(lane, basen + A * b);
            }
        }
        if (FUSE2) {  // the factor of this butterfly was fetched one iteration ago; fetch the next one now so that
                      // its L2 latency hides behind this butterfly
#pragma unroll
            for (int d = 0; d < R; ++d) f[d] = QEB_VPF ? fnext[d] : mid.fetch(lane, base + d);
            const int jn = j + jstep;
            if (QEB_VPF && jn < NJ) {
                const int blkn = jn / A;
                const int basen = blkn * L + (jn - blkn * A);
#pragma unroll
                for (int d = 0; d < R; ++d) fnext[d] = mid.fetch(lane, basen + d);
            }
        }
        if (!PF) {
#pragma unroll
            for (int b = 0; b < R; ++b) v[b] = src.load(lane, base + A * b);
        }
        if (DIT && A > 1 && a != 0) {  // a == 0: every twiddle is 1 (warp-uniform test)
            const int st = a * NB;
#pragma unroll
            for (int b = 1; b < R; ++b) v[b] = cmul(v[b], stw(tw, st * b, sgn));
        }
        Bfly<R>::run(v, sgn);
        if (FUSE2) {
#pragma unroll
            for (int d = 0; d < R; ++d) v[d] = scal(v[d], f[d]);
            Bfly<R>::run(v, -sgn);
        }
        if (!DIT && !FUSE2 && A > 1 && a != 0) {
            const int st = a * NB;
#pragma unroll
            for (int d = 1; d < R; ++d) v[d] = cmul(v[d], stw(tw, st * d, sgn));
        }
#pragma unroll
        for (int d = 0; d < R; ++d) dst.store(lane, base + A * d, v[d]);
    }
}

// Spinor form of the fused middle pass: last DIF pass (A = 1, no twiddles) of BOTH components of a spinor, the 2x2
// pointwise mix (mix.apply(lane, pos, up, dn)), first DIT pass of both. 2*R complex values live per thread.
template <class PL, class SrcU, class SrcD, class Mix>
QEB_HD void spass_spin(int warp, int nwarps, int lane, int nl, double sgn, SrcU& su, SrcD& sd, Mix& mix)
{
    constexpr int R = PL::r(PL::NP - 1);
    constexpr int NJ = PL::N / R;
    if (lane >= nl) return;
    for (int j = warp; j < NJ; j += nwarps) {
        const int base = j * R;
        cplx u[R], d[R];
        double pv[R][4];  // the 2x2 potential of these R points: issued first, so that the global latency hides behind
                          // the shared-memory loads and the two DIF butterflies
#pragma unroll
        for (int k = 0; k < R; ++k) mix.fetch(lane, base + k, pv[k]);
#pragma unroll
        for (int b = 0; b < R; ++b) {
            u[b] = su.load(lane, base + b);
            d[b] = sd.load(lane, base + b);
        }
        Bfly<R>::run(u, sgn);
        Bfly<R>::run(d, sgn);
#pragma unroll
        for (int k = 0; k < R; ++k) mix.apply(pv[k], u[k], d[k]);
        Bfly<R>::run(u, -sgn);
        Bfly<R>::run(d, -sgn);
#pragma unroll
        for (int k = 0; k < R; ++k) {
            su.store(lane, base + k, u[k]);
            sd.store(lane, base + k, d[k]);
        }
    }
}

#if defined(__CUDACC__)
// Spinor middle pass, half-warp layout: lanes 0-15 carry the up component of 16 lines, lanes 16-31 the down component
// of the same lines, so one ordinary tile holds the spinor and the 2x2 mix is a register exchange with lane^16 between
// the last DIF and the first DIT butterfly. Every lane of the warp takes part (idle lines compute on zeros).
// Device only: the lane exchange has no serial emulation (tests: GPU parity at the AUSURF112 grid).
template <class PL, class Src, class Mix>
__device__ __forceinline__ void spass_spinx(int warp, int nwarps, int lane, double sgn, Src& src, Mix& mix)
{
    constexpr int R = PL::r(PL::NP - 1);
    constexpr int NJ = PL::N / R;
    const double s = lane < 16 ? 1.0 : -1.0;
    for (int j = warp; j < NJ; j += nwarps) {
        const int base = j * R;
        cplx v[R];
        double pv[R][4];
#pragma unroll
        for (int k = 0; k < R; ++k) mix.fetch(lane, base + k, pv[k]);
#pragma unroll
        for (int b = 0; b < R; ++b) v[b] = src.load(lane, base + b);
        Bfly<R>::run(v, sgn);
#pragma unroll
        for (int k = 0; k < R; ++k) {
            cplx o;
            o.x = __shfl_xor_sync(0xffffffffu, v[k].x, 16);
            o.y = __shfl_xor_sync(0xffffffffu, v[k].y, 16);
            // up:   psi_up (v1+v4) + psi_dn (v2 - i v3);  down: psi_dn (v1-v4) + psi_up (v2 + i v3)
            const double a = pv[k][0] + s * pv[k][3];
            const double sv3 = s * pv[k][2];
            v[k] = mk(v[k].x * a + (o.x * pv[k][1] + o.y * sv3), v[k].y * a + (o.y * pv[k][1] - o.x * sv3));
        }
        Bfly<R>::run(v, -sgn);
#pragma unroll
        for (int k = 0; k < R; ++k) src.store(lane, base + k, v[k]);
    }
}
#endif

// run pass number `p` (runtime, warp-uniform) of plan PL: expands to a switch over the NP passes
template <class PL, bool DIT, bool FUSE2, int P = 0, bool PF = false, int SUB = 1>
struct SPassSel {
    template <class Src, class Dst, class Mid>
    QEB_HD static void run(int p, int warp, int nwarps, int lane, int nl, const cplx* tw, double sgn, Src& src, Dst& dst,
                           Mid& mid)
    {
        if (p == P)
            spass<PL, P, DIT, FUSE2, PF, SUB>(warp, nwarps, lane, nl, tw, sgn, src, dst, mid);
        else if constexpr (P + 1 < PL::NP)
            SPassSel<PL, DIT, FUSE2, P + 1, PF, SUB>::run(p, warp, nwarps, lane, nl, tw, sgn, src, dst, mid);
    }
};

// the static plans that are instantiated (BASELINE grids 25, 27, 64, 125, 135, 144, 200 + common small sizes);
// DIF order, identical to fft_plan.h::preferred_factors (large in-register radices: at most one
// shared-memory exchange per transform)
using SP15 = SPlan<15, 15>;
using SP25 = SPlan<25, 25>;
using SP27 = SPlan<27, 27>;
using SP64 = SPlan<64, 8, 8>;
using SP125 = SPlan<125, 25, 5>;
using SP135 = SPlan<135, 15, 9>;
using SP144 = SPlan<144, 12, 12>;
using SP200 = SPlan<200, 25, 8>;
// low-register alternatives (radices <= 8: about half the registers per thread, twice the resident warps, one more
// shared-memory exchange); selected with QEB_XPLAN / QEB_YPLAN / QEB_ZPLAN
using SP200b = SPlan<200, 8, 5, 5>;
using SP200c = SPlan<200, 20, 10>;
using SP200d = SPlan<200, 10, 20>;
using SP125b = SPlan<125, 5, 5, 5>;
using SP64b = SPlan<64, 4, 4, 4>;

}  // namespace qeb
