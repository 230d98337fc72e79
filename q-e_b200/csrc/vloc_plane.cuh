// Fused plane pass of the V_loc path: for grids whose (x,y) plane fits in shared memory, ONE kernel replaces
// P2 (y inverse), P3 (x inverse * V * x forward) and P4 (y forward) of vloc_kernels.cuh, so the G2 intermediate
// (nxw*nr2*nr3 complex per box, written and read twice) never exists:
//
//   G1[box][z][stick] --> plane(z) in shared memory: y DIF(+) | x DIF(+) * V(r)/N * x DIT(-) | y DIT(-) --> G1
//
// (reference: the cft_2xy pair around the V(r) multiply, FFTXlib/src/fft_scalar.FFTW3.f90:197-365 and
//  PW/src/vloc_psi_acc.f90:197-229; the slab path does the same per z plane, fft_parallel_2d.f90:36-204.)
//
// One CTA owns plane z of `bpc` consecutive boxes. The plane lives as P[ypos*ldp + x] (ldp odd: the x passes,
// whose lanes run along ypos, then hit distinct 16-byte bank groups); y stays in DIF position order between
// the y passes, so no digit-reversal copy exists (V is pre-permuted in x and y: pl_v[z][xpos][ypos]).
// The stick row of the next box is fetched with cp.async while the current plane is transformed.
// Work mapping as in fft_static.cuh: lane = line (active x column in the y passes, ypos row in the x passes),
// warps stride over (line group, butterfly index).
#pragma once
#include "fft_static.cuh"
#include "vloc_kernels.cuh"

namespace qeb {

struct PlaneSm {      // tables the plane kernel stages in shared memory (the CPU emulation points them at host arrays)
    const cplx* twx;
    const cplx* twy;
    const int* ymapT;  // [nr2*nxw] stick id at (y, xa) or -1
    const int* xmap;   // [nr1]     x -> xa or -1
    const int* xact;   // [nxw]     xa -> x
};

struct PlYCompactSrc {  // first y DIF pass: natural y -> stick of this active column (pruned)
    const cplx* st;
    const int* ymapT;   // at column xa0 of the lane group
    int nxw;
    QEB_HD cplx load(int lane, int y) const
    {
        const int s = ymapT[y * nxw + lane];
        return s >= 0 ? st[s] : mk(0.0, 0.0);
    }
};
struct PlYCompactDst {  // last y DIT pass: natural y -> stick (others dropped)
    cplx* st;
    const int* ymapT;
    int nxw;
    QEB_HD void store(int lane, int y, cplx v) const
    {
        const int s = ymapT[y * nxw + lane];
        if (s >= 0) st[s] = v;
    }
};
struct PlCol {  // y passes on the plane: this lane's active column, rows = DIF positions
    cplx* col;
    int ldp;
    QEB_HD cplx load(int, int pos) const { return col[(size_t)pos * ldp]; }
    QEB_HD void store(int, int pos, cplx v) const { col[(size_t)pos * ldp] = v; }
};
struct PlRowNat {  // first / last x pass: natural x, inactive columns read as zero and are not stored
    cplx* row;
    const int* xmap;
    QEB_HD cplx load(int, int x) const { return xmap[x] >= 0 ? row[x] : mk(0.0, 0.0); }
    QEB_HD void store(int, int x, cplx v) const
    {
        if (xmap[x] >= 0) row[x] = v;
    }
};
struct PlRow {  // x passes in position order
    cplx* row;
    QEB_HD cplx load(int, int pos) const { return row[pos]; }
    QEB_HD void store(int, int pos, cplx v) const { row[pos] = v; }
};
struct PlVMid {  // V(r)/N of this plane, [xpos][ypos]
    const double* vz;  // at (z, 0, y0 + lane)
    int nr2;
    QEB_HD double fetch(int, int xpos) const
    {
#if defined(__CUDA_ARCH__)
        return __ldg(vz + (size_t)xpos * nr2);
#else
        return vz[(size_t)xpos * nr2];
#endif
    }
};

template <class PLX, class PLY>
QEB_HD int plane_nphase()
{
    return 2 * PLY::NP + 2 * PLX::NP - 1;
}

// One barrier-separated phase of the plane kernel for plane z: stin / stout are the compact stick rows
// (nst complex) read by phase 0 and written by the last phase.
template <class PLX, class PLY>
QEB_HD void plane_phase(int ph, int tid, int nthr, const VlocTables& t, int z, cplx* P, const cplx* stin, cplx* stout,
                        const PlaneSm& sm)
{
    constexpr int npy = PLY::NP, npx = PLX::NP;
    const int nwt = nthr / QEB_TL, w = tid / QEB_TL, lane = tid % QEB_TL;
    if (w >= nwt) return;
    const int ldp = t.pl_ldp;
    NoMid nm;
    const bool ypass = ph < npy || ph >= npy + 2 * npx - 1;
    const int nlines = ypass ? t.nxw : t.nr2;
    const int ng = (nlines + QEB_TL - 1) / QEB_TL;
    const int gstep = nwt < ng ? nwt : ng;  // line groups worked on concurrently
    const int g0 = w % gstep;
    const int wsub = w / gstep;
    const int nws = (nwt - g0 + gstep - 1) / gstep;  // warps that share this warp's groups
    for (int g = g0; g < ng; g += gstep) {
        const int l0 = g * QEB_TL;
        const int nl = (nlines - l0) < QEB_TL ? (nlines - l0) : QEB_TL;
        const int ln = lane < nl ? lane : 0;  // idle lanes keep valid addresses (they return inside spass)
        if (ypass) {
            PlCol c{P + sm.xact[l0 + ln], ldp};
            if (ph < npy) {  // y inverse, DIF pass p
                const int p = ph;
                if (p == 0) {
                    PlYCompactSrc cs{stin, sm.ymapT + l0, t.nxw};
                    SPassSel<PLY, false, false>::run(p, wsub, nws, lane, nl, sm.twy, +1.0, cs, c, nm);
                } else {
                    SPassSel<PLY, false, false>::run(p, wsub, nws, lane, nl, sm.twy, +1.0, c, c, nm);
                }
            } else {  // y forward, DIT pass p
                const int p = npy - 1 - (ph - (npy + 2 * npx - 1));
                if (p == 0) {
                    PlYCompactDst cd{stout, sm.ymapT + l0, t.nxw};
                    SPassSel<PLY, true, false>::run(p, wsub, nws, lane, nl, sm.twy, -1.0, c, cd, nm);
                } else {
                    SPassSel<PLY, true, false>::run(p, wsub, nws, lane, nl, sm.twy, -1.0, c, c, nm);
                }
            }
        } else {  // x phases: same list as xfus_fast_phase
            const int q = ph - npy;
            cplx* row = P + (size_t)(l0 + ln) * ldp;
            PlRowNat gnat{row, sm.xmap};
            PlRow wk{row};
            PlVMid vm{t.pl_v + ((size_t)z * t.nr1) * t.nr2 + l0 + ln, t.nr2};
            if (q < npx - 1) {
                if (q == 0)
                    SPassSel<PLX, false, false>::run(q, wsub, nws, lane, nl, sm.twx, +1.0, gnat, wk, nm);
                else
                    SPassSel<PLX, false, false>::run(q, wsub, nws, lane, nl, sm.twx, +1.0, wk, wk, nm);
            } else if (q == npx - 1) {
                if (npx == 1)
                    SPassSel<PLX, false, true>::run(npx - 1, wsub, nws, lane, nl, sm.twx, +1.0, gnat, gnat, vm);
                else
                    SPassSel<PLX, false, true>::run(npx - 1, wsub, nws, lane, nl, sm.twx, +1.0, wk, wk, vm);
            } else {
                const int p = 2 * npx - 2 - q;
                if (p == 0)
                    SPassSel<PLX, true, false>::run(p, wsub, nws, lane, nl, sm.twx, -1.0, wk, gnat, nm);
                else
                    SPassSel<PLX, true, false>::run(p, wsub, nws, lane, nl, sm.twx, -1.0, wk, wk, nm);
            }
        }
    }
}

// shared-memory layout of the plane kernel (bytes); nstp = sticks rounded up to 4
QEB_HD size_t plane_smem_bytes(int nr1, int nr2, int nst, int nxw, int ldp)
{
    const size_t nstp = ((size_t)nst + 3) & ~(size_t)3;
    const size_t c16 = (size_t)nr2 * ldp + 2 * nstp + nr1 + nr2;
    const size_t i4 = (size_t)nr2 * nxw + nr1 + nxw;
    return c16 * 16 + ((i4 * 4 + 15) & ~(size_t)15) + 16;
}

}  // namespace qeb
