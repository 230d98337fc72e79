// Full-box in-place 3-D FFT (the fwfft/invfft entry points of
// FFTXlib/src/fft_interfaces.f90:19-80 on a serial descriptor; replaces
// cfft3d / cfft3d_gpu, fft_scalar.FFTW3.f90:379-484, fft_scalar.cuFFT.f90:523-654).
// Box layout: x fastest, f[x + nr1*(y + nr2*z)]. Three pencil kernels:
//   z: lines along z, tile = T consecutive columns (coalesced along the column index)
//   y: lines along y, tile = T consecutive x at fixed z
//   x: lines along x, tile = T consecutive (y,z) rows, transposed through padded smem
// sign > 0: exp(+i) unscaled; sign < 0: exp(-i), scaled by 1/(nr1 nr2 nr3) in the x kernel.
#pragma once
#include "fft_core.cuh"

namespace qeb {

struct FullFftParams {
    int nr1, nr2, nr3;
    FftPlan px, py, pz;
    cplx* f;
    double sgn, scale;
    int TZ, TYx, TXr;  // tile sizes
    FastDiv dTZ, dTZl, dTY, dTYl, dTX, dTXl;
};

struct StridedGlobal {  // element pos of line l at base[pos_map(pos)*estride + l]
    cplx* base;
    size_t estride;
    const int* rev;  // nullptr: identity
    QEB_HD cplx load(int line, int pos) const { return base[(size_t)(rev ? rev[pos] : pos) * estride + line]; }
    QEB_HD void store(int line, int pos, cplx v) const { base[(size_t)(rev ? rev[pos] : pos) * estride + line] = v; }
};

// generic strided pencil: DIF, natural in (from global), digit-reversal resolved at the final store
QEB_HD void pencil_phase(int ph, int tid, int nthr, const FftPlan& pl, double sgn, cplx* gbase, size_t estride, int nl,
                         const FastDiv& dnl, int T, cplx* smem)
{
    const int np = pl.npass;
    const PassDesc pd = pass_of(pl, ph);
    SmemTile s{smem, T};
    StridedGlobal gin{gbase, estride, nullptr}, gout{gbase, estride, pl.rev};
    NoMid nm;
    if (np == 1)
        fft_pass<false, false>(tid, nthr, nl, dnl, pd, pl.tw, sgn, gin, gout, nm);
    else if (ph == 0)
        fft_pass<false, false>(tid, nthr, nl, dnl, pd, pl.tw, sgn, gin, s, nm);
    else if (ph == np - 1)
        fft_pass<false, false>(tid, nthr, nl, dnl, pd, pl.tw, sgn, s, gout, nm);
    else
        fft_pass<false, false>(tid, nthr, nl, dnl, pd, pl.tw, sgn, s, s, nm);
}

QEB_HD void fullz_phase(int ph, int tid, int nthr, const FullFftParams& p, int tile, int box, cplx* smem)
{
    const int ncol = p.nr1 * p.nr2;
    const int c0 = tile * p.TZ;
    const int nl = (ncol - c0) < p.TZ ? (ncol - c0) : p.TZ;
    cplx* g = p.f + (size_t)box * ncol * p.nr3 + c0;
    pencil_phase(ph, tid, nthr, p.pz, p.sgn, g, (size_t)ncol, nl, nl == p.TZ ? p.dTZ : p.dTZl, p.TZ, smem);
}

QEB_HD void fully_phase(int ph, int tid, int nthr, const FullFftParams& p, int tile, int z, int box, cplx* smem)
{
    const int x0 = tile * p.TYx;
    const int nl = (p.nr1 - x0) < p.TYx ? (p.nr1 - x0) : p.TYx;
    cplx* g = p.f + ((size_t)box * p.nr3 + z) * p.nr1 * p.nr2 + x0;
    pencil_phase(ph, tid, nthr, p.py, p.sgn, g, (size_t)p.nr1, nl, nl == p.TYx ? p.dTY : p.dTYl, p.TYx, smem);
}

// x: phase 0 load (transposing), phases 1..np passes in smem, phase np+1 store (natural order, scaled)
QEB_HD int fullx_nphase(const FullFftParams& p) { return p.px.npass + 2; }
QEB_HD void fullx_phase(int ph, int tid, int nthr, const FullFftParams& p, int tile, int box, cplx* smem)
{
    const int nrows = p.nr2 * p.nr3;
    const int r0 = tile * p.TXr;
    const int nl = (nrows - r0) < p.TXr ? (nrows - r0) : p.TXr;
    const int ld = p.TXr + 1;
    cplx* g = p.f + ((size_t)box * nrows + r0) * p.nr1;
    const int np = p.px.npass;
    if (ph == 0) {
        const int tot = nl * p.nr1;
        for (int i = tid; i < tot; i += nthr) {
            const int row = i / p.nr1, x = i - row * p.nr1;
            smem[x * ld + row] = g[i];
        }
        return;
    }
    if (ph <= np) {
        const PassDesc pd = pass_of(p.px, ph - 1);
        SmemTile s{smem, ld};
        NoMid nm;
        fft_pass<false, false>(tid, nthr, nl, nl == p.TXr ? p.dTX : p.dTXl, pd, p.px.tw, p.sgn, s, s, nm);
        return;
    }
    // smem position pos holds X[rev[pos]]: read through the inverse map so that global writes stay coalesced
    const int tot = nl * p.nr1;
    for (int i = tid; i < tot; i += nthr) {
        const int row = i / p.nr1, pos = i - row * p.nr1;
        const cplx v = smem[pos * ld + row];
        g[(size_t)row * p.nr1 + p.px.rev[pos]] = scal(v, p.scale);
    }
}

}  // namespace qeb
