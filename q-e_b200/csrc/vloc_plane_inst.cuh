// Plane kernel (vloc_plane.cuh): kernel template + launcher, instantiated once per (x plan, y plan) pair in
// vloc_plane_<n1>_<n2>.cu so that the instantiations compile in parallel.
#pragma once
#include <cuda_runtime.h>
#include <stdlib.h>
#include "vloc.h"
#include "vloc_plane.cuh"

namespace qeb {


// named barriers over one line group between passes of the same direction: measured SLOWER than CTA barriers
// (V_loc 26.97 -> 27.86 us/band at AUSURF112); off unless QEB_USE_PLANE_GROUPBAR is defined
#ifdef QEB_USE_PLANE_GROUPBAR
constexpr bool QEB_PLANE_GROUPBAR = true;
#else
constexpr bool QEB_PLANE_GROUPBAR = false;
#endif

QEB_D void pl_cp_async16(void* smem_dst, const void* gsrc)
{
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gsrc));
}
QEB_D void pl_cp_async_wait_all() { asm volatile("cp.async.wait_all;\n" ::: "memory"); }

// MAXT: threads per CTA the kernel is compiled for (512: 128 registers, radix-25 butterflies; 1024: 64 registers,
// radices <= 8, twice the warps)
template <class PLX, class PLY, int MAXT>
__global__ void __launch_bounds__(MAXT) k_vloc_plane_f(const __grid_constant__ VlocTables t, int nboxes, int bpc)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    cplx* P = reinterpret_cast<cplx*>(smem_raw);
    const int ldp = t.pl_ldp, nst = t.nst, nxw = t.nxw;
    const int nstp = (nst + 3) & ~3;
    cplx* stin = P + (size_t)t.nr2 * ldp;
    cplx* stout = stin + nstp;
    cplx* twx = stout + nstp;
    cplx* twy = twx + PLX::N;
    int* ymapT = reinterpret_cast<int*>(twy + PLY::N);
    int* xmapS = ymapT + (size_t)t.nr2 * nxw;
    int* xactS = xmapS + t.nr1;
    const int tid = threadIdx.x, nthr = blockDim.x;
    const int z = blockIdx.x;
    const int b0 = blockIdx.y * bpc;
    const int b1 = (b0 + bpc) < nboxes ? (b0 + bpc) : nboxes;
    // stick row of the first box: in flight while the tables are staged
    {
        const cplx* row = t.g1 + (size_t)b0 * t.g1_box + (size_t)z * t.g1_zs;
        for (int i = tid; i < nst; i += nthr) pl_cp_async16(stin + i, row + i);
    }
    for (int i = tid; i < PLX::N; i += nthr) twx[i] = t.px.tw[i];
    for (int i = tid; i < PLY::N; i += nthr) twy[i] = t.py.tw[i];
    for (int i = tid; i < t.nr2 * nxw; i += nthr) ymapT[i] = t.pl_ymapT[i];
    for (int i = tid; i < t.nr1; i += nthr) xmapS[i] = t.xmap[i];
    for (int i = tid; i < nxw; i += nthr) xactS[i] = t.pl_xact[i];
    const PlaneSm sm{twx, twy, ymapT, xmapS, xactS};
    constexpr int nph = 2 * PLY::NP + 2 * PLX::NP - 1;
    for (int b = b0; b < b1; ++b) {
        pl_cp_async_wait_all();
        __syncthreads();
#pragma unroll 1
        for (int ph = 0; ph < nph; ++ph) {
            plane_phase<PLX, PLY>(ph, tid, nthr, t, z, P, stin, stout, sm);
            // Between two passes of the same direction only the warps of one line group exchange data (group g = warp
            // % ng owns lines 32g..32g+31 in both passes): a named barrier over that group replaces the CTA barrier.
            // ph == 0 keeps the CTA barrier: the prefetch below overwrites stin, which every group reads in phase 0.
            const bool ycur = ph < PLY::NP || ph >= PLY::NP + 2 * PLX::NP - 1;
            const bool ynxt = (ph + 1) < PLY::NP || (ph + 1) >= PLY::NP + 2 * PLX::NP - 1;
            const int ng = ((ycur ? nxw : t.nr2) + QEB_TL - 1) / QEB_TL;
            const int nwt = nthr / QEB_TL;
            if (QEB_PLANE_GROUPBAR && ph > 0 && ph + 1 < nph && ycur == ynxt && ng > 1 && ng <= 15 && nwt % ng == 0) {
                const int g = (tid / QEB_TL) % ng;
                asm volatile("bar.sync %0, %1;" ::"r"(1 + g), "r"((nwt / ng) * QEB_TL) : "memory");
            } else {
                __syncthreads();
            }
            if (ph == 0 && b + 1 < b1) {  // stin is consumed: fetch the next box's row behind the remaining phases
                const cplx* row = t.g1 + (size_t)(b + 1) * t.g1_box + (size_t)z * t.g1_zs;
                for (int i = tid; i < nst; i += nthr) pl_cp_async16(stin + i, row + i);
            }
        }
        cplx* orow = t.g1 + (size_t)b * t.g1_box + (size_t)z * t.g1_zs;
        for (int i = tid; i < nst; i += nthr) orow[i] = stout[i];
    }
}

template <class PLX, class PLY, int MAXT = 512>
struct PlaneLaunch {
    static cudaError_t attrs(int s);
    // false: this pair does not match the tables' plans
    static bool run(const VlocTables& t, int nboxes, int num_sms, size_t max_smem, cudaStream_t st);
};

#ifdef QEB_PLANE_INSTANTIATE
static inline int pl_env_int(const char* name, int dflt)
{
    const char* e = getenv(name);
    return e ? atoi(e) : dflt;
}
template <class PL>
static int pl_max_nj()
{
    int m = 1;
    for (int p = 0; p < PL::NP; ++p) m = (PL::N / PL::r(p)) > m ? (PL::N / PL::r(p)) : m;
    return m;
}
template <class PLX, class PLY, int MAXT>
cudaError_t PlaneLaunch<PLX, PLY, MAXT>::attrs(int s)
{
    return cudaFuncSetAttribute(k_vloc_plane_f<PLX, PLY, MAXT>, cudaFuncAttributeMaxDynamicSharedMemorySize, s);
}
template <class PLX, class PLY, int MAXT>
bool PlaneLaunch<PLX, PLY, MAXT>::run(const VlocTables& t, int nboxes, int num_sms, size_t max_smem, cudaStream_t st)
{
    if (!PLX::matches(t.px) || !PLY::matches(t.py)) return false;
    const size_t smem = plane_smem_bytes(t.nr1, t.nr2, t.nst, t.nxw, t.pl_ldp);
    const int ngx = (t.nr2 + QEB_TL - 1) / QEB_TL, ngy = (t.nxw + QEB_TL - 1) / QEB_TL;
    int tasks = ngx * pl_max_nj<PLX>();
    if (ngy * pl_max_nj<PLY>() > tasks) tasks = ngy * pl_max_nj<PLY>();
    int nw = tasks < MAXT / 32 ? tasks : MAXT / 32;
    nw = pl_env_int("QEB_NWP", nw);
    if (nw < 1) nw = 1;
    if (nw * 32 > MAXT) nw = MAXT / 32;
    // boxes per CTA: the fewest "plane slots" per SM over the whole launch (ties: more boxes per CTA = more prefetch)
    int per_sm = (int)(max_smem / smem);
    if (per_sm > 2048 / (nw * 32)) per_sm = 2048 / (nw * 32);
    if (per_sm > 32) per_sm = 32;
    if (per_sm < 1) per_sm = 1;
    const long slots = (long)num_sms * per_sm;
    // one box per CTA has no next stick row to prefetch: measured 8 % slower per plane (profiles/r1n_sweep.log), so it
    // only wins when it saves more than that in wave quantisation
    int bpc = 1;
    double best = -1.0;
    for (int c = 1; c <= 8 && c <= nboxes; ++c) {
        const long ctas = (long)t.nr3 * ((nboxes + c - 1) / c);
        const double cost = (double)(((ctas + slots - 1) / slots) * c) * (c == 1 ? 1.08 : 1.0);
        if (best < 0.0 || cost <= best) {
            best = cost;
            bpc = c;
        }
    }
    bpc = pl_env_int("QEB_PLANE_BPC", bpc);
    if (bpc < 1) bpc = 1;
    const dim3 grid(t.nr3, (nboxes + bpc - 1) / bpc);
    k_vloc_plane_f<PLX, PLY, MAXT><<<grid, nw * 32, smem, st>>>(t, nboxes, bpc);
    return true;
}
#endif

}  // namespace qeb
