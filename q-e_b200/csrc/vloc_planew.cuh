// Fused plane pass, warp-local form ("planew"): the same transform as vloc_plane.cuh
//
//   G1[box][z][stick] --> plane(z) in shared memory: y DIF(+) | x DIF(+) * V(r)/N * x DIT(-) | y DIT(-) --> G1
//
// (the cft_2xy pair around the V(r) multiply, FFTXlib/src/fft_scalar.FFTW3.f90:197-365, PW/src/vloc_psi_acc.f90:197-229)
// with a different work mapping. vloc_plane.cuh gives one line to a lane and lets the warps of the CTA stride over
// the butterfly index, so every radix pass ends in a CTA barrier (7 per plane at 125 x 64) and the 16 resident warps
// alternate in lock-step between a shared-memory burst and an FP64 burst. Here a line belongs to a TEAM of T lanes
// of ONE warp (T = gcd of the butterfly counts of the plan's passes: 8 for 64 = 8*8, 5 for 125 = 25*5), a warp owns
// 32/T lines for ALL passes of a direction, and consecutive passes are separated by __syncwarp() only. CTA barriers
// remain at the two direction changes (y -> x, x -> y) and at the stick-row hand-over: 4 per plane. Warps drift
// apart inside a direction, so the shared-memory traffic of one warp overlaps the butterflies of another.
//
// Plane layout: element (ypos, x) at P[ypos*ldp + ysh(ypos) + x], ysh(ypos) = ypos / R_last(y) when that radix is even
// (else 0). With ldp odd and ldp = T_x (mod 8) the 16-byte accesses of a quarter warp fall into distinct bank groups:
//   y passes  (lanes: team member fastest)  stride between members = ldp (odd), or R_last*ldp + 1 in the A == 1 pass
//   x passes  members take consecutive units of a row; consecutive rows are T_x units apart (mod 8)
// V is read as in vloc_plane.cuh (pl_v[z][xpos][ypos], pre-permuted in x and y). Executed on the CPU by tests/emul.
#pragma once
#include "vloc_plane.cuh"

namespace qeb {

template <class PLY>
struct PlwShift {  // ysh(ypos)
    static constexpr int RL = PLY::r(PLY::NP - 1);
    static constexpr bool ON = (RL % 2) == 0;
    QEB_HD static int of(int ypos) { return ON ? ypos / RL : 0; }
    QEB_HD static int max_shift(int nr2) { return ON ? (nr2 - 1) / RL : 0; }
};

// row stride of the planew layout: odd, >= nr1 + max shift, = T_x (mod 8) when T_x is odd
template <class PLX, class PLY>
QEB_HD int planew_ldp(int nr1, int nr2)
{
    int ld = nr1 + PlwShift<PLY>::max_shift(nr2);
    if ((ld & 1) == 0) ++ld;
    constexpr int TX = STeam<PLX>::T;
    if (TX & 1)
        while ((ld & 7) != (TX & 7)) ld += 2;
    return ld;
}

template <class PLY>
struct PlwCol {  // y passes: this thread's active column, rows = DIF positions, shifted layout
    cplx* col;   // P + x
    int ldp;
    QEB_HD cplx load(int, int pos) const { return col[(size_t)pos * ldp + PlwShift<PLY>::of(pos)]; }
    QEB_HD void store(int, int pos, cplx v) const { col[(size_t)pos * ldp + PlwShift<PLY>::of(pos)] = v; }
};

template <class PLX, class PLY>
QEB_HD int planew_nsteps()
{
    return 2 * PLY::NP + 2 * PLX::NP - 1;
}
// does a CTA barrier (not just __syncwarp) follow step s? (direction changes and the last step)
template <class PLX, class PLY>
QEB_HD bool planew_cta_barrier_after(int s)
{
    return s == PLY::NP - 1 || s == PLY::NP + 2 * PLX::NP - 2 || s == planew_nsteps<PLX, PLY>() - 1;
}

// One step (= one radix pass of one direction) of plane z. Same step list as plane_phase. ldp = planew_ldp().
template <class PLX, class PLY>
QEB_HD void planew_step(int s, int tid, int nthr, const VlocTables& t, int z, cplx* P, const cplx* stin, cplx* stout,
                        const PlaneSm& sm, int ldp)
{
    constexpr int npy = PLY::NP, npx = PLX::NP;
    const int nwt = nthr / 32, w = tid / 32, lane = tid % 32;
    if (w >= nwt) return;
    NoMid nm;
    const bool ypass = s < npy || s >= npy + 2 * npx - 1;
    if (ypass) {
        constexpr int T = STeam<PLY>::T, LPW = STeam<PLY>::LPW;
        const int nlines = t.nxw;
        const int ntask = (nlines + LPW - 1) / LPW;
        const int li = lane / T, tm = lane - li * T;
        if (li >= LPW) return;
        for (int task = w; task < ntask; task += nwt) {
            const int l = task * LPW + li;
            if (l >= nlines) continue;
            PlwCol<PLY> c{P + sm.xact[l], ldp};
            if (s < npy) {  // y inverse, DIF pass s
                if (s == 0) {
                    PlYCompactSrc cs{stin, sm.ymapT + l, t.nxw};
                    SPassSelT<PLY, false, false>::run(s, tm, sm.twy, +1.0, cs, c, nm);
                } else {
                    SPassSelT<PLY, false, false>::run(s, tm, sm.twy, +1.0, c, c, nm);
                }
            } else {  // y forward, DIT pass p
                const int p = npy - 1 - (s - (npy + 2 * npx - 1));
                if (p == 0) {
                    PlYCompactDst cd{stout, sm.ymapT + l, t.nxw};
                    SPassSelT<PLY, true, false>::run(p, tm, sm.twy, -1.0, c, cd, nm);
                } else {
                    SPassSelT<PLY, true, false>::run(p, tm, sm.twy, -1.0, c, c, nm);
                }
            }
        }
    } else {
        constexpr int T = STeam<PLX>::T, LPW = STeam<PLX>::LPW;
        const int nlines = t.nr2;
        const int ntask = (nlines + LPW - 1) / LPW;
        const int li = lane / T, tm = lane - li * T;
        if (li >= LPW) return;
        const int q = s - npy;
        for (int task = w; task < ntask; task += nwt) {
            const int l = task * LPW + li;  // ypos row
            if (l >= nlines) continue;
            cplx* row = P + (size_t)l * ldp + PlwShift<PLY>::of(l);
            PlRowNat gnat{row, sm.xmap};
            PlRow wk{row};
            PlVMid vm{t.pl_v + ((size_t)z * t.nr1) * t.nr2 + l, t.nr2};
            if (q < npx - 1) {
                if (q == 0)
                    SPassSelT<PLX, false, false>::run(q, tm, sm.twx, +1.0, gnat, wk, nm);
                else
                    SPassSelT<PLX, false, false>::run(q, tm, sm.twx, +1.0, wk, wk, nm);
            } else if (q == npx - 1) {
                if (npx == 1)
                    SPassSelT<PLX, false, true>::run(npx - 1, tm, sm.twx, +1.0, gnat, gnat, vm);
                else
                    SPassSelT<PLX, false, true>::run(npx - 1, tm, sm.twx, +1.0, wk, wk, vm);
            } else {
                const int p = 2 * npx - 2 - q;
                if (p == 0)
                    SPassSelT<PLX, true, false>::run(p, tm, sm.twx, -1.0, wk, gnat, nm);
                else
                    SPassSelT<PLX, true, false>::run(p, tm, sm.twx, -1.0, wk, wk, nm);
            }
        }
    }
}

// shared-memory bytes of the planew kernel
QEB_HD size_t planew_smem_bytes(int nr1, int nr2, int nst, int nxw, int ldp, int max_shift)
{
    const size_t nstp = ((size_t)nst + 3) & ~(size_t)3;
    const size_t c16 = (size_t)nr2 * ldp + max_shift + 2 * nstp + nr1 + nr2;
    const size_t i4 = (size_t)nr2 * nxw + nr1 + nxw;
    return c16 * 16 + ((i4 * 4 + 15) & ~(size_t)15) + 16;
}

}  // namespace qeb
