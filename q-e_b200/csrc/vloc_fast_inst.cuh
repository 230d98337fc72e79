// Fast V_loc kernels (compile-time FFT plan): kernel templates + launchers, instantiated once per static
// plan in vloc_fast_<N>.cu so that the instantiations compile in parallel.
#pragma once
#include <cuda_runtime.h>
#include <stdlib.h>
#include "vloc.h"
#include "vloc_fast.cuh"

namespace qeb {

#ifndef QEB_FAST_MAXTHREADS
#define QEB_FAST_MAXTHREADS 512
#endif
#define QEB_SMEM extern __shared__ __align__(16) unsigned char smem_raw[]; cplx* smem = reinterpret_cast<cplx*>(smem_raw)

// ---- fast variants: compile-time plan of the pass direction, warp-mapped work ----
template <class PL>
__global__ void __launch_bounds__(QEB_FAST_MAXTHREADS) k_vloc_zinv_f(const __grid_constant__ VlocTables t, const __grid_constant__ VlocIO io)
{
    QEB_SMEM;
    cplx* twS = smem + ((size_t)t.nr3 * t.TZ);
    int* tabS = reinterpret_cast<int*>(twS + PL::N);
    for (int i = threadIdx.x; i < PL::N; i += blockDim.x) twS[i] = t.pz.tw[i];
    for (int i = threadIdx.x; i < PL::N; i += blockDim.x) tabS[i] = t.pz.rev[i];
    for (int i = threadIdx.x; i < t.TZ; i += blockDim.x) tabS[PL::N + i] = t.grp_stick[(size_t)blockIdx.x * t.TZ + i];
    const FastTabs ft{twS, nullptr, tabS, tabS + PL::N};
    __syncthreads();
    for (int ph = 0; ph < 2 + PL::NP; ++ph) {
        zinv_fast_phase<PL>(ph, threadIdx.x, blockDim.x, t, io, blockIdx.x, blockIdx.y, smem, ft);
        __syncthreads();
    }
}
template <class PL>
__global__ void __launch_bounds__(QEB_FAST_MAXTHREADS) k_vloc_zfwd_f(const __grid_constant__ VlocTables t, const __grid_constant__ VlocIO io)
{
    QEB_SMEM;
    cplx* twS = smem + ((size_t)t.nr3 * t.TZ);
    int* tabS = reinterpret_cast<int*>(twS + PL::N);
    for (int i = threadIdx.x; i < PL::N; i += blockDim.x) twS[i] = t.pz.tw[i];
    for (int i = threadIdx.x; i < PL::N; i += blockDim.x) tabS[i] = t.pz.rev[i];
    for (int i = threadIdx.x; i < t.TZ; i += blockDim.x) tabS[PL::N + i] = t.grp_stick[(size_t)blockIdx.x * t.TZ + i];
    const FastTabs ft{twS, nullptr, tabS, tabS + PL::N};
    __syncthreads();
    for (int ph = 0; ph < 1 + PL::NP; ++ph) {
        zfwd_fast_phase<PL>(ph, threadIdx.x, blockDim.x, t, io, blockIdx.x, blockIdx.y, smem, ft);
        __syncthreads();
    }
}
template <class PL>
__global__ void __launch_bounds__(QEB_FAST_MAXTHREADS) k_vloc_yinv_f(const __grid_constant__ VlocTables t)
{
    QEB_SMEM;
    cplx* twS = smem + ((size_t)t.nsmax * (t.TY + 1) + (size_t)t.nr2 * t.TY);
    int* tabS = reinterpret_cast<int*>(twS + PL::N);
    for (int i = threadIdx.x; i < PL::N; i += blockDim.x) twS[i] = t.py.tw[i];
    for (int i = threadIdx.x; i < PL::N; i += blockDim.x) {
        tabS[i] = t.py.rev[i];
        tabS[PL::N + i] = t.ymap[(size_t)blockIdx.y * t.nr2 + i];
    }
    const FastTabs ft{twS, tabS + PL::N, tabS, nullptr};
    __syncthreads();
    for (int ph = 0; ph < 1 + PL::NP; ++ph) {
        yinv_fast_phase<PL>(ph, threadIdx.x, blockDim.x, t, blockIdx.x, blockIdx.y, blockIdx.z, smem, ft);
        __syncthreads();
    }
}
template <class PL>
__global__ void __launch_bounds__(QEB_FAST_MAXTHREADS) k_vloc_yfwd_f(const __grid_constant__ VlocTables t)
{
    QEB_SMEM;
    cplx* twS = smem + ((size_t)t.nsmax * (t.TY + 1) + (size_t)t.nr2 * t.TY);
    int* tabS = reinterpret_cast<int*>(twS + PL::N);
    for (int i = threadIdx.x; i < PL::N; i += blockDim.x) twS[i] = t.py.tw[i];
    for (int i = threadIdx.x; i < PL::N; i += blockDim.x) {
        tabS[i] = t.py.rev[i];
        tabS[PL::N + i] = t.ymap[(size_t)blockIdx.y * t.nr2 + i];
    }
    const FastTabs ft{twS, tabS + PL::N, tabS, nullptr};
    __syncthreads();
    for (int ph = 0; ph < 1 + PL::NP; ++ph) {
        yfwd_fast_phase<PL>(ph, threadIdx.x, blockDim.x, t, blockIdx.x, blockIdx.y, blockIdx.z, smem, ft);
        __syncthreads();
    }
}
template <class PL>
__global__ void __launch_bounds__(QEB_FAST_MAXTHREADS) k_vloc_xfus_f(const __grid_constant__ VlocTables t)
{
    QEB_SMEM;
    cplx* twS = smem + ((size_t)t.nr1 * t.TX);
    int* tabS = reinterpret_cast<int*>(twS + PL::N);
    for (int i = threadIdx.x; i < PL::N; i += blockDim.x) twS[i] = t.px.tw[i];
    for (int i = threadIdx.x; i < PL::N; i += blockDim.x) tabS[i] = t.xmap[i];
    const FastTabs ft{twS, tabS, nullptr, nullptr};
    __syncthreads();
    for (int ph = 0; ph < 2 * PL::NP - 1; ++ph) {
        xfus_fast_phase<PL>(ph, threadIdx.x, blockDim.x, t, blockIdx.x, blockIdx.y, smem, ft);
        __syncthreads();
    }
}


// warps per CTA for a plan: the count in [lo,hi] that wastes the fewest warp-rounds over the passes
template <class PL>
static int pick_warps(int lo, int hi)
{
    int best = lo;
    double beff = -1.0;
    for (int nw = lo; nw <= hi; ++nw) {
        long used = 0, tot = 0;
        for (int p = 0; p < PL::NP; ++p) {
            const int nj = PL::N / PL::r(p);
            used += nj;
            tot += (long)((nj + nw - 1) / nw) * nw;
        }
        const double eff = (double)used / (double)tot;
        if (eff > beff - 1e-9) {  // ties: more warps
            beff = eff;
            best = nw;
        }
    }
    return best;
}

static inline int env_int(const char* name, int dflt)
{
    const char* e = getenv(name);
    return e ? atoi(e) : dflt;
}

template <class PL>
struct FastLaunch {
    static void attrs(int s);
    static void z(bool inv, const VlocTables& t, const VlocIO& io, const VlocLaunchCfg& cfg, int nboxes, cudaStream_t st);
    static void y(bool inv, const VlocTables& t, const VlocLaunchCfg& cfg, int nboxes, cudaStream_t st);
    static void x(const VlocTables& t, const VlocLaunchCfg& cfg, int nboxes, cudaStream_t st);
};

#ifdef QEB_FAST_INSTANTIATE
template <class PL>
void FastLaunch<PL>::attrs(int s)
{
    cudaFuncSetAttribute(k_vloc_zinv_f<PL>, cudaFuncAttributeMaxDynamicSharedMemorySize, s);
    cudaFuncSetAttribute(k_vloc_zfwd_f<PL>, cudaFuncAttributeMaxDynamicSharedMemorySize, s);
    cudaFuncSetAttribute(k_vloc_yinv_f<PL>, cudaFuncAttributeMaxDynamicSharedMemorySize, s);
    cudaFuncSetAttribute(k_vloc_yfwd_f<PL>, cudaFuncAttributeMaxDynamicSharedMemorySize, s);
    cudaFuncSetAttribute(k_vloc_xfus_f<PL>, cudaFuncAttributeMaxDynamicSharedMemorySize, s);
}
template <class PL>
void FastLaunch<PL>::z(bool inv, const VlocTables& t, const VlocIO& io, const VlocLaunchCfg& cfg, int nboxes, cudaStream_t st)
{
    const dim3 grid(t.ngrp, nboxes);
    const int nw = env_int("QEB_NWZ", pick_warps<PL>(4, 10));
    if (inv)
        k_vloc_zinv_f<PL><<<grid, nw * 32, cfg.smem_z + PL::N * 20 + 4 * QEB_TL + 16, st>>>(t, io);
    else
        k_vloc_zfwd_f<PL><<<grid, nw * 32, cfg.smem_z + PL::N * 20 + 4 * QEB_TL + 16, st>>>(t, io);
}
template <class PL>
void FastLaunch<PL>::y(bool inv, const VlocTables& t, const VlocLaunchCfg& cfg, int nboxes, cudaStream_t st)
{
    const dim3 grid(t.nty, t.nxw, nboxes);
    const int nw = env_int("QEB_NWY", pick_warps<PL>(4, 10));
    if (inv)
        k_vloc_yinv_f<PL><<<grid, nw * 32, cfg.smem_y + PL::N * 24 + 16, st>>>(t);
    else
        k_vloc_yfwd_f<PL><<<grid, nw * 32, cfg.smem_y + PL::N * 24 + 16, st>>>(t);
}
template <class PL>
void FastLaunch<PL>::x(const VlocTables& t, const VlocLaunchCfg& cfg, int nboxes, cudaStream_t st)
{
    const dim3 grid(t.ntx, nboxes);
    const int nw = env_int("QEB_NWX", pick_warps<PL>(4, 10));
    k_vloc_xfus_f<PL><<<grid, nw * 32, cfg.smem_x + PL::N * 20 + 16, st>>>(t);
}
#endif

}  // namespace qeb
