// planew kernel (vloc_planew.cuh): kernel template + launcher, instantiated once per (x plan, y plan) pair in
// vloc_planew_<n1>_<n2>.cu so that the instantiations compile in parallel.
#pragma once
#include <cuda_runtime.h>
#include <stdlib.h>
#include "vloc.h"
#include "vloc_planew.cuh"

namespace qeb {

QEB_D void plw_cp_async16(void* smem_dst, const void* gsrc)
{
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gsrc));
}
QEB_D void plw_cp_async_wait_all() { asm volatile("cp.async.wait_all;\n" ::: "memory"); }

template <class PLX, class PLY, int MAXT>
__global__ void __launch_bounds__(MAXT) k_vloc_planew_f(const __grid_constant__ VlocTables t, int nboxes, int bpc, int ldp)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    cplx* P = reinterpret_cast<cplx*>(smem_raw);
    const int nst = t.nst, nxw = t.nxw;
    const int nstp = (nst + 3) & ~3;
    cplx* stin = P + (size_t)t.nr2 * ldp + PlwShift<PLY>::max_shift(t.nr2);
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
    {  // stick row of the first box: in flight while the tables are staged
        const cplx* row = t.g1 + (size_t)b0 * t.g1_box + (size_t)z * t.g1_zs;
        for (int i = tid; i < nst; i += nthr) plw_cp_async16(stin + i, row + i);
    }
    for (int i = tid; i < PLX::N; i += nthr) twx[i] = t.px.tw[i];
    for (int i = tid; i < PLY::N; i += nthr) twy[i] = t.py.tw[i];
    for (int i = tid; i < t.nr2 * nxw; i += nthr) ymapT[i] = t.pl_ymapT[i];
    for (int i = tid; i < t.nr1; i += nthr) xmapS[i] = t.xmap[i];
    for (int i = tid; i < nxw; i += nthr) xactS[i] = t.pl_xact[i];
    const PlaneSm sm{twx, twy, ymapT, xmapS, xactS};
    constexpr int ns = 2 * PLY::NP + 2 * PLX::NP - 1;
    for (int b = b0; b < b1; ++b) {
        plw_cp_async_wait_all();
        __syncthreads();  // stin of this box landed; stout of the previous box has been copied out
#pragma unroll 1
        for (int s = 0; s < ns; ++s) {
            planew_step<PLX, PLY>(s, tid, nthr, t, z, P, stin, stout, sm, ldp);
            if (planew_cta_barrier_after<PLX, PLY>(s))
                __syncthreads();
            else
                __syncwarp();
            if (s == PLY::NP - 1 && b + 1 < b1) {  // every warp is past the first y pass: stin is free for the next box
                const cplx* row = t.g1 + (size_t)(b + 1) * t.g1_box + (size_t)z * t.g1_zs;
                for (int i = tid; i < nst; i += nthr) plw_cp_async16(stin + i, row + i);
            }
        }
        cplx* orow = t.g1 + (size_t)b * t.g1_box + (size_t)z * t.g1_zs;
        for (int i = tid; i < nst; i += nthr) orow[i] = stout[i];
    }
}

template <class PLX, class PLY, int MAXT = 512>
struct PlaneWLaunch {
    static cudaError_t attrs(int s);
    static size_t smem(const VlocTables& t);
    // false: this pair does not match the tables' plans, or the plane does not fit
    static bool run(const VlocTables& t, int nboxes, int num_sms, size_t max_smem, cudaStream_t st);
};

#ifdef QEB_PLANEW_INSTANTIATE
static inline int plw_env_int(const char* name, int dflt)
{
    const char* e = getenv(name);
    return e ? atoi(e) : dflt;
}
template <class PLX, class PLY, int MAXT>
cudaError_t PlaneWLaunch<PLX, PLY, MAXT>::attrs(int s)
{
    return cudaFuncSetAttribute(k_vloc_planew_f<PLX, PLY, MAXT>, cudaFuncAttributeMaxDynamicSharedMemorySize, s);
}
template <class PLX, class PLY, int MAXT>
size_t PlaneWLaunch<PLX, PLY, MAXT>::smem(const VlocTables& t)
{
    return planew_smem_bytes(t.nr1, t.nr2, t.nst, t.nxw, planew_ldp<PLX, PLY>(t.nr1, t.nr2), PlwShift<PLY>::max_shift(t.nr2));
}
template <class PLX, class PLY, int MAXT>
bool PlaneWLaunch<PLX, PLY, MAXT>::run(const VlocTables& t, int nboxes, int num_sms, size_t max_smem, cudaStream_t st)
{
    if (!PLX::matches(t.px) || !PLY::matches(t.py)) return false;
    const int ldp = planew_ldp<PLX, PLY>(t.nr1, t.nr2);
    const size_t sm = smem(t);
    if (sm > max_smem) return false;
    // warps: enough for the direction with more warp-tasks (a task = the 32/T lines one warp carries through a direction)
    const int ty = (t.nxw + STeam<PLY>::LPW - 1) / STeam<PLY>::LPW, tx = (t.nr2 + STeam<PLX>::LPW - 1) / STeam<PLX>::LPW;
    int nw = ty > tx ? ty : tx;
    if (nw > MAXT / 32) nw = MAXT / 32;
    nw = plw_env_int("QEB_NWP", nw);
    if (nw < 1) nw = 1;
    if (nw * 32 > MAXT) nw = MAXT / 32;
    int per_sm = (int)(max_smem / sm);
    if (per_sm > 2048 / (nw * 32)) per_sm = 2048 / (nw * 32);
    if (per_sm > 32) per_sm = 32;
    if (per_sm < 1) per_sm = 1;
    const long slots = (long)num_sms * per_sm;
    int bpc = 1;
    double best = -1.0;
    for (int c = 1; c <= 8 && c <= nboxes; ++c) {  // as PlaneLaunch::run: fewest plane slots per SM, prefetch bonus
        const long ctas = (long)t.nr3 * ((nboxes + c - 1) / c);
        const double cost = (double)(((ctas + slots - 1) / slots) * c) * (c == 1 ? 1.08 : 1.0);
        if (best < 0.0 || cost <= best) {
            best = cost;
            bpc = c;
        }
    }
    bpc = plw_env_int("QEB_PLANE_BPC", bpc);
    if (bpc < 1) bpc = 1;
    const dim3 grid(t.nr3, (nboxes + bpc - 1) / bpc);
    k_vloc_planew_f<PLX, PLY, MAXT><<<grid, nw * 32, sm, st>>>(t, nboxes, bpc, ldp);
    return true;
}
#endif

}  // namespace qeb
