// Radix butterflies and the generic in-place radix pass used by every FFT
// kernel of the V_loc path (replaces the 1-D transforms the reference takes
// from FFTW3 / cuFFT: FFTXlib/src/fft_scalar.FFTW3.f90:498-694,
// fft_scalar.cuFFT.f90:523-654).
//
// Convention: bfly<R>(v, sgn) computes X[k] = sum_n v[n] exp(sgn*2*pi*i*n*k/R),
// natural order in and out, in registers. sgn=+1 is the reference's G->R
// transform (invfft), sgn=-1 its R->G transform (fwfft, scaling applied
// elsewhere).
#pragma once
#include "common.cuh"
#include "tw_const.h"

namespace qeb {

template <int R>
struct Bfly;

template <>
struct Bfly<2> {
    static QEB_HD void run(cplx* v, double)
    {
        cplx a = v[0] + v[1], b = v[0] - v[1];
        v[0] = a;
        v[1] = b;
    }
};

template <>
struct Bfly<3> {
    static QEB_HD void run(cplx* v, double sgn)
    {
        const double s3 = 0.86602540378443865 * sgn;  // sin(2pi/3)
        cplx t = v[1] + v[2];
        cplx d = v[1] - v[2];
        cplx m = mk(v[0].x - 0.5 * t.x, v[0].y - 0.5 * t.y);
        v[0] = v[0] + t;
        // i*s3*d = (-s3*d.y, s3*d.x)
        v[1] = mk(m.x - s3 * d.y, m.y + s3 * d.x);
        v[2] = mk(m.x + s3 * d.y, m.y - s3 * d.x);
    }
};

template <>
struct Bfly<4> {
    static QEB_HD void run(cplx* v, double sgn)
    {
        cplx a = v[0] + v[2], b = v[0] - v[2];
        cplx c = v[1] + v[3], d = v[1] - v[3];
        cplx id = muli(d, sgn);
        v[0] = a + c;
        v[2] = a - c;
        v[1] = b + id;
        v[3] = b - id;
    }
};

template <>
struct Bfly<5> {
    static QEB_HD void run(cplx* v, double sgn)
    {
        const double c1 = 0.30901699437494742, c2 = -0.80901699437494742;
        const double s1 = 0.95105651629515357 * sgn, s2 = 0.58778525229247313 * sgn;
        cplx t1 = v[1] + v[4], t2 = v[2] + v[3];
        cplx t3 = v[1] - v[4], t4 = v[2] - v[3];
        cplx m1 = mk(v[0].x + c1 * t1.x + c2 * t2.x, v[0].y + c1 * t1.y + c2 * t2.y);
        cplx m2 = mk(v[0].x + c2 * t1.x + c1 * t2.x, v[0].y + c2 * t1.y + c1 * t2.y);
        cplx u1 = mk(s1 * t3.x + s2 * t4.x, s1 * t3.y + s2 * t4.y);
        cplx u2 = mk(s2 * t3.x - s1 * t4.x, s2 * t3.y - s1 * t4.y);
        v[0] = v[0] + t1 + t2;
        v[1] = mk(m1.x - u1.y, m1.y + u1.x);
        v[4] = mk(m1.x + u1.y, m1.y - u1.x);
        v[2] = mk(m2.x - u2.y, m2.y + u2.x);
        v[3] = mk(m2.x + u2.y, m2.y - u2.x);
    }
};

template <>
struct Bfly<8> {
    static QEB_HD void run(cplx* v, double sgn)
    {
        const double h = 0.70710678118654752;
        cplx a0 = v[0] + v[4], b0 = v[0] - v[4];
        cplx a1 = v[1] + v[5], b1 = v[1] - v[5];
        cplx a2 = v[2] + v[6], b2 = v[2] - v[6];
        cplx a3 = v[3] + v[7], b3 = v[3] - v[7];
        // odd branch twiddles w8^j, w8 = (1 + sgn*i)/sqrt2
        cplx t1 = mk((b1.x - sgn * b1.y) * h, (sgn * b1.x + b1.y) * h);
        cplx t2 = muli(b2, sgn);
        cplx t3 = mk((-b3.x - sgn * b3.y) * h, (sgn * b3.x - b3.y) * h);
        // even: DFT4(a0..a3) -> X[0],X[2],X[4],X[6]
        {
            cplx p = a0 + a2, q = a0 - a2, r = a1 + a3, s = muli(a1 - a3, sgn);
            v[0] = p + r;
            v[4] = p - r;
            v[2] = q + s;
            v[6] = q - s;
        }
        // odd: DFT4(b0,t1,t2,t3) -> X[1],X[3],X[5],X[7]
        {
            cplx p = b0 + t2, q = b0 - t2, r = t1 + t3, s = muli(t1 - t3, sgn);
            v[1] = p + r;
            v[5] = p - r;
            v[3] = q + s;
            v[7] = q - s;
        }
    }
};

// Composite radix R = P*Q in registers (DIF split, constant twiddles):
// n = a + Q*b, k = P*c + d:  X[P*c+d] = sum_a w_Q^{ac} [ w_R^{ad} sum_b v[a+Q*b] w_P^{bd} ].
template <int R, int P, int Q>
struct BflyComp {
    static QEB_HD void run(cplx* v, double sgn)
    {
        cplx y[R];  // y[a*P + d]
#pragma unroll
        for (int a = 0; a < Q; ++a) {
            cplx t[P];
#pragma unroll
            for (int b = 0; b < P; ++b) t[b] = v[a + Q * b];
            Bfly<P>::run(t, sgn);
#pragma unroll
            for (int d = 0; d < P; ++d) {
                const int e = (a * d) % R;
                if (e == 0) {
                    y[a * P + d] = t[d];
                } else if (4 * e == R) {
                    y[a * P + d] = muli(t[d], sgn);
                } else if (2 * e == R) {
                    y[a * P + d] = mk(-t[d].x, -t[d].y);
                } else if (4 * e == 3 * R) {
                    y[a * P + d] = muli(t[d], -sgn);
                } else {
                    const double wr = twc_re(R, e), wi = sgn * twc_im(R, e);
                    y[a * P + d] = mk(t[d].x * wr - t[d].y * wi, t[d].x * wi + t[d].y * wr);
                }
            }
        }
#pragma unroll
        for (int d = 0; d < P; ++d) {
            cplx t[Q];
#pragma unroll
            for (int a = 0; a < Q; ++a) t[a] = y[a * P + d];
            Bfly<Q>::run(t, sgn);
#pragma unroll
            for (int c = 0; c < Q; ++c) v[P * c + d] = t[c];
        }
    }
};

template <> struct Bfly<6> { static QEB_HD void run(cplx* v, double s) { BflyComp<6, 2, 3>::run(v, s); } };
template <> struct Bfly<9> { static QEB_HD void run(cplx* v, double s) { BflyComp<9, 3, 3>::run(v, s); } };
template <> struct Bfly<10> { static QEB_HD void run(cplx* v, double s) { BflyComp<10, 2, 5>::run(v, s); } };
template <> struct Bfly<12> { static QEB_HD void run(cplx* v, double s) { BflyComp<12, 4, 3>::run(v, s); } };
template <> struct Bfly<15> { static QEB_HD void run(cplx* v, double s) { BflyComp<15, 3, 5>::run(v, s); } };
template <> struct Bfly<16> { static QEB_HD void run(cplx* v, double s) { BflyComp<16, 4, 4>::run(v, s); } };
template <> struct Bfly<20> { static QEB_HD void run(cplx* v, double s) { BflyComp<20, 4, 5>::run(v, s); } };
template <> struct Bfly<25> { static QEB_HD void run(cplx* v, double s) { BflyComp<25, 5, 5>::run(v, s); } };
template <> struct Bfly<27> { static QEB_HD void run(cplx* v, double s) { BflyComp<27, 3, 9>::run(v, s); } };

// twiddle exp(sgn*2*pi*i*j/N) from the +sign table
QEB_HD cplx twid(const cplx* tw, int j, double sgn)
{
#if defined(__CUDA_ARCH__)
    const double2 t = __ldg(reinterpret_cast<const double2*>(tw) + j);
    return mk(t.x, sgn * t.y);
#else
    return mk(tw[j].x, sgn * tw[j].y);
#endif
}

// Descriptor of one pass of one plan, flattened for the kernels.
struct PassDesc {
    int r, L, A, nb;  // radix, sub-length, stride A = L/r, nb = N/L
    FastDiv divA;
};

QEB_HD PassDesc pass_of(const FftPlan& pl, int p)
{
    PassDesc d;
    d.r = pl.r[p];
    d.L = pl.L[p];
    d.A = d.L / d.r;
    d.nb = pl.n / d.L;
    d.divA = pl.divA[p];
    return d;
}

// One radix-R pass over a tile of nl lines.  Work item w -> (line = w % nl,
// j = w / nl); j -> (blk, a); positions base + A*b with base = blk*L + a.
//   DIF: butterfly, then twiddle w_L^{a d} on output d.
//   DIT: twiddle w_L^{a b} on input b, then butterfly.
// MID (optional) is applied between two butterflies when FUSE2 is set: the
// pass is then "last DIF pass (A=1) -> pointwise op -> first DIT pass" of the
// fused x-direction kernel (both passes have A = 1 and no twiddles).
template <int R, bool DIT, bool FUSE2, class Src, class Dst, class Mid>
QEB_HD void fft_pass_r(int tid, int nthr, int nl, const FastDiv& divNl, const PassDesc& pd, const cplx* tw,
                       double sgn, Src& src, Dst& dst, Mid& mid)
{
    const int nitems = nl * (pd.L / R) * pd.nb;
    for (int w = tid; w < nitems; w += nthr) {
        const uint32_t j = fd_div(divNl, (uint32_t)w);
        const int line = w - (int)j * nl;
        const uint32_t blk = fd_div(pd.divA, j);
        const int a = (int)j - (int)blk * pd.A;
        const int base = (int)blk * pd.L + a;
        cplx v[R];
#pragma unroll
        for (int b = 0; b < R; ++b) v[b] = src.load(line, base + pd.A * b);
        if (DIT && pd.A > 1) {
            const int st = a * pd.nb;
#pragma unroll
            for (int b = 1; b < R; ++b) v[b] = cmul(v[b], twid(tw, st * b, sgn));
        }
        Bfly<R>::run(v, sgn);
        if (FUSE2) {
#pragma unroll
            for (int d = 0; d < R; ++d) v[d] = mid.apply(line, base + d, v[d]);
            Bfly<R>::run(v, -sgn);
        }
        if (!DIT && !FUSE2 && pd.A > 1) {
            const int st = a * pd.nb;
#pragma unroll
            for (int d = 1; d < R; ++d) v[d] = cmul(v[d], twid(tw, st * d, sgn));
        }
#pragma unroll
        for (int d = 0; d < R; ++d) dst.store(line, base + pd.A * d, v[d]);
    }
}

struct NoMid {
    QEB_HD cplx apply(int, int, cplx v) const { return v; }
    QEB_HD double fetch(int, int) const { return 1.0; }
};

// Radix dispatch (warp-uniform switch).
template <bool DIT, bool FUSE2, class Src, class Dst, class Mid>
QEB_HD void fft_pass(int tid, int nthr, int nl, const FastDiv& divNl, const PassDesc& pd, const cplx* tw, double sgn,
                     Src& src, Dst& dst, Mid& mid)
{
    switch (pd.r) {
    case 2: fft_pass_r<2, DIT, FUSE2>(tid, nthr, nl, divNl, pd, tw, sgn, src, dst, mid); break;
    case 3: fft_pass_r<3, DIT, FUSE2>(tid, nthr, nl, divNl, pd, tw, sgn, src, dst, mid); break;
    case 4: fft_pass_r<4, DIT, FUSE2>(tid, nthr, nl, divNl, pd, tw, sgn, src, dst, mid); break;
    case 5: fft_pass_r<5, DIT, FUSE2>(tid, nthr, nl, divNl, pd, tw, sgn, src, dst, mid); break;
    case 6: fft_pass_r<6, DIT, FUSE2>(tid, nthr, nl, divNl, pd, tw, sgn, src, dst, mid); break;
    case 8: fft_pass_r<8, DIT, FUSE2>(tid, nthr, nl, divNl, pd, tw, sgn, src, dst, mid); break;
    case 9: fft_pass_r<9, DIT, FUSE2>(tid, nthr, nl, divNl, pd, tw, sgn, src, dst, mid); break;
    case 10: fft_pass_r<10, DIT, FUSE2>(tid, nthr, nl, divNl, pd, tw, sgn, src, dst, mid); break;
    case 12: fft_pass_r<12, DIT, FUSE2>(tid, nthr, nl, divNl, pd, tw, sgn, src, dst, mid); break;
    case 15: fft_pass_r<15, DIT, FUSE2>(tid, nthr, nl, divNl, pd, tw, sgn, src, dst, mid); break;
    case 16: fft_pass_r<16, DIT, FUSE2>(tid, nthr, nl, divNl, pd, tw, sgn, src, dst, mid); break;
    case 20: fft_pass_r<20, DIT, FUSE2>(tid, nthr, nl, divNl, pd, tw, sgn, src, dst, mid); break;
    case 25: fft_pass_r<25, DIT, FUSE2>(tid, nthr, nl, divNl, pd, tw, sgn, src, dst, mid); break;
    case 27: fft_pass_r<27, DIT, FUSE2>(tid, nthr, nl, divNl, pd, tw, sgn, src, dst, mid); break;
    default: break;
    }
}

// shared-memory tile accessor: S[pos*ld + line]
struct SmemTile {
    cplx* s;
    int ld;
    QEB_HD cplx load(int line, int pos) const { return s[pos * ld + line]; }
    QEB_HD void store(int line, int pos, cplx v) const { s[pos * ld + line] = v; }
};

}  // namespace qeb
