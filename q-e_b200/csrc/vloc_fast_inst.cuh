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
// OCC: resident CTAs per SM the register budget is compiled for (1: 128 registers with up to 512 threads; 4: 160 threads
// and 102 registers, so that four 51 KB tiles fit an SM)
template <class PL, int OCC>
__global__ void __launch_bounds__(OCC > 1 ? 160 : QEB_FAST_MAXTHREADS, OCC) k_vloc_zinv_f(const __grid_constant__ VlocTables t, const __grid_constant__ VlocIO io)
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
        zinv_fast_phase<PL, 1>(ph, threadIdx.x, blockDim.x, t, io, blockIdx.x, blockIdx.y, smem, ft);
        __syncthreads();
    }
}
template <class PL, bool STAGE, int OCC>
__global__ void __launch_bounds__(OCC > 1 ? 160 : (STAGE ? 256 : QEB_FAST_MAXTHREADS), OCC > 1 ? OCC : (STAGE ? 2 : 1))
    k_vloc_zfwd_f(const __grid_constant__ VlocTables t, const __grid_constant__ VlocIO io)
{
    QEB_SMEM;
    cplx* twS = smem + ((size_t)t.nr3 * t.TZ);
    int* tabS = reinterpret_cast<int*>(twS + PL::N);
    int* irevS = tabS + PL::N + QEB_TL;
    for (int i = threadIdx.x; i < PL::N; i += blockDim.x) {
        const int r = t.pz.rev[i];
        tabS[i] = r;
        if (STAGE) irevS[r] = i;
    }
    for (int i = threadIdx.x; i < PL::N; i += blockDim.x) twS[i] = t.pz.tw[i];
    for (int i = threadIdx.x; i < t.TZ; i += blockDim.x) tabS[PL::N + i] = t.grp_stick[(size_t)blockIdx.x * t.TZ + i];
    const FastTabs ft{twS, nullptr, tabS, tabS + PL::N};
    __syncthreads();
    if (STAGE) {
        zfwd_stage(threadIdx.x, blockDim.x, t, blockIdx.x, blockIdx.y, smem, irevS, ft.sticks);
        asm volatile("cp.async.wait_all;\n" ::: "memory");
        __syncthreads();
    }
    for (int ph = 0; ph < 1 + PL::NP; ++ph) {
        zfwd_fast_phase<PL, 1, STAGE>(ph, threadIdx.x, blockDim.x, t, io, blockIdx.x, blockIdx.y, smem, ft);
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

// spinor x pass: block = (line tile, band); two tiles
template <class PL>
__global__ void __launch_bounds__(QEB_FAST_MAXTHREADS) k_vloc_xspin_f(const __grid_constant__ VlocTables t)
{
    QEB_SMEM;
    cplx* twS = smem + 2 * ((size_t)t.nr1 * t.TX);
    int* tabS = reinterpret_cast<int*>(twS + PL::N);
    for (int i = threadIdx.x; i < PL::N; i += blockDim.x) twS[i] = t.px.tw[i];
    for (int i = threadIdx.x; i < PL::N; i += blockDim.x) tabS[i] = t.xmap[i];
    const FastTabs ft{twS, tabS, nullptr, nullptr};
    __syncthreads();
    for (int ph = 0; ph < 2 * PL::NP - 1; ++ph) {
        // band = fast grid index: the CTAs that share a line tile (the same slice of the four potentials) run together
        xspin_fast_phase<PL>(ph, threadIdx.x, blockDim.x, t, blockIdx.y, blockIdx.x, smem, ft);
        __syncthreads();
    }
}

// spinor x pass, half-warp layout (fft_static.cuh::spass_spinx): block = (band, tile of 16 lines); ONE tile of 32 lanes
// = 16 lines x 2 components, so the shared-memory footprint and the occupancy are those of the scalar x pass
struct G2SpinX {  // lane -> (component = lane / 16, line = lane % 16); natural x, pruned; idle lines read zero
    cplx* g2u;    // box of the up component; the down component is the next box
    size_t g2_box;
    const int* xmap;
    int npos, p0, nl;
    __device__ __forceinline__ cplx load(int lane, int x) const
    {
        const int line = lane & 15, xa = xmap[x];
        if (xa < 0 || line >= nl) return mk(0.0, 0.0);
        return g2u[(size_t)(lane >> 4) * g2_box + (size_t)xa * npos + p0 + line];
    }
    __device__ __forceinline__ void store(int lane, int x, cplx v) const
    {
        const int line = lane & 15, xa = xmap[x];
        if (xa >= 0 && line < nl) g2u[(size_t)(lane >> 4) * g2_box + (size_t)xa * npos + p0 + line] = v;
    }
};
struct SpinMixX {
    const double* vnc;
    size_t vs;
    int npos, p0, nl;
    __device__ __forceinline__ void fetch(int lane, int xpos, double v[4]) const
    {
        const int line = (lane & 15) < nl ? (lane & 15) : 0;
        const size_t iv = (size_t)xpos * npos + p0 + line;
        v[0] = __ldg(vnc + iv), v[1] = __ldg(vnc + vs + iv), v[2] = __ldg(vnc + 2 * vs + iv), v[3] = __ldg(vnc + 3 * vs + iv);
    }
};
template <class PL>
__global__ void __launch_bounds__(QEB_FAST_MAXTHREADS) k_vloc_xspinx_f(const __grid_constant__ VlocTables t)
{
    QEB_SMEM;
    constexpr int np = PL::NP;
    cplx* twS = smem + ((size_t)t.nr1 * QEB_TL);
    int* tabS = reinterpret_cast<int*>(twS + PL::N);
    for (int i = threadIdx.x; i < PL::N; i += blockDim.x) twS[i] = t.px.tw[i];
    for (int i = threadIdx.x; i < PL::N; i += blockDim.x) tabS[i] = t.xmap[i];
    __syncthreads();
    const int pair = blockIdx.x, p0 = blockIdx.y * 16;
    const int nl = (t.npos - p0) < 16 ? (t.npos - p0) : 16;
    const int warp = threadIdx.x / QEB_TL, lane = threadIdx.x % QEB_TL, nwarps = blockDim.x / QEB_TL;
    LaneTile w{smem, QEB_TL};
    G2SpinX g{t.g2 + (size_t)(2 * pair) * t.g2_box, t.g2_box, tabS, t.npos, p0, nl};
    SpinMixX mix{t.vnc, (size_t)t.nr1 * t.npos, t.npos, p0, nl};
    NoMid nm;
    for (int ph = 0; ph < 2 * np - 1; ++ph) {
        if (ph < np - 1) {
            if (ph == 0)
                SPassSel<PL, false, false>::run(ph, warp, nwarps, lane, QEB_TL, twS, +1.0, g, w, nm);
            else
                SPassSel<PL, false, false>::run(ph, warp, nwarps, lane, QEB_TL, twS, +1.0, w, w, nm);
        } else if (ph == np - 1) {
            if (np == 1)
                spass_spinx<PL>(warp, nwarps, lane, +1.0, g, mix);
            else
                spass_spinx<PL>(warp, nwarps, lane, +1.0, w, mix);
        } else {
            const int p = 2 * np - 2 - ph;
            if (p == 0)
                SPassSel<PL, true, false>::run(p, warp, nwarps, lane, QEB_TL, twS, -1.0, w, g, nm);
            else
                SPassSel<PL, true, false>::run(p, warp, nwarps, lane, QEB_TL, twS, -1.0, w, w, nm);
        }
        __syncthreads();
    }
}

// density accumulation (sum_band): block = (line tile, box chunk); the CTA x-transforms its boxes of the batch one after
// the other and accumulates w|psi(r)|^2 in shared memory, then adds the tile to rho (atomically when the batch is
// split over several chunks)
template <class PL>
__global__ void __launch_bounds__(QEB_FAST_MAXTHREADS) k_vloc_xacc_f(const __grid_constant__ VlocTables t, const __grid_constant__ VlocIO io,
                                                                   const double* __restrict__ w, int nboxes, int bpc,
                                                                   double* __restrict__ rho)
{
    QEB_SMEM;
    double* racc = reinterpret_cast<double*>(smem + (size_t)t.nr1 * t.TX);
    // (the accumulator is padded to an even count: nr1 * TX is odd for an odd grid with an odd tile, e.g. the one-plane
    // rank of a 15^3 grid distributed over 8 ranks, and the complex twiddles behind it need 16-byte alignment)
    cplx* twS = reinterpret_cast<cplx*>(racc + (((size_t)t.nr1 * t.TX + 1) & ~(size_t)1));
    int* tabS = reinterpret_cast<int*>(twS + PL::N);
    int* irev = tabS + PL::N;
    for (int i = threadIdx.x; i < PL::N; i += blockDim.x) {
        twS[i] = t.px.tw[i];
        tabS[i] = t.xmap[i];
        irev[t.px.rev[i]] = i;
    }
    xacc_zero_phase(threadIdx.x, blockDim.x, t, racc);
    const FastTabs ft{twS, tabS, nullptr, nullptr};
    __syncthreads();
    const int b0 = blockIdx.y * bpc;
    const int b1 = (b0 + bpc) < nboxes ? (b0 + bpc) : nboxes;
    for (int box = b0; box < b1; ++box)
        for (int ph = 0; ph < PL::NP; ++ph) {
            xacc_fast_phase<PL>(ph, threadIdx.x, blockDim.x, t, io, w, blockIdx.x, box, smem, racc, ft);
            __syncthreads();
        }
    xacc_flush_phase(threadIdx.x, blockDim.x, t, blockIdx.x, irev, racc, rho, gridDim.y > 1);
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
    static void xacc(const VlocTables& t, const VlocIO& io, const VlocLaunchCfg& cfg, int nboxes, const double* w,
                     double* rho, cudaStream_t st);
    static void xspin(const VlocTables& t, const VlocLaunchCfg& cfg, int nbands, cudaStream_t st);
};

#ifdef QEB_FAST_INSTANTIATE
template <class PL>
void FastLaunch<PL>::attrs(int s)
{
    cudaFuncSetAttribute(k_vloc_zinv_f<PL, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, s);
    cudaFuncSetAttribute(k_vloc_zinv_f<PL, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, s);
    cudaFuncSetAttribute(k_vloc_zfwd_f<PL, false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, s);
    cudaFuncSetAttribute(k_vloc_zfwd_f<PL, true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, s);
    cudaFuncSetAttribute(k_vloc_zfwd_f<PL, true, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, s);
    cudaFuncSetAttribute(k_vloc_yinv_f<PL>, cudaFuncAttributeMaxDynamicSharedMemorySize, s);
    cudaFuncSetAttribute(k_vloc_yfwd_f<PL>, cudaFuncAttributeMaxDynamicSharedMemorySize, s);
    cudaFuncSetAttribute(k_vloc_xfus_f<PL>, cudaFuncAttributeMaxDynamicSharedMemorySize, s);
    cudaFuncSetAttribute(k_vloc_xacc_f<PL>, cudaFuncAttributeMaxDynamicSharedMemorySize, s);
    cudaFuncSetAttribute(k_vloc_xspin_f<PL>, cudaFuncAttributeMaxDynamicSharedMemorySize, s);
    cudaFuncSetAttribute(k_vloc_xspinx_f<PL>, cudaFuncAttributeMaxDynamicSharedMemorySize, s);
}
template <class PL>
void FastLaunch<PL>::z(bool inv, const VlocTables& t, const VlocIO& io, const VlocLaunchCfg& cfg, int nboxes, cudaStream_t st)
{
    const dim3 grid(t.ngrp, nboxes);
    const bool stage = !inv && env_int("QEB_ZSTAGE", 1) != 0;
    const size_t smem = cfg.smem_z + PL::N * 24 + 4 * QEB_TL + 16;
    // QEB_ZOCC=4: the 96-register build with four resident CTAs per SM; measured slower (spills: 22.9 vs 20.8 us/band at
    // AUSURF112, profiles/r2j_sweep.log), off by default
    const bool occ4 = env_int("QEB_ZOCC", 1) == 4 && 4 * (smem + 1024) <= cfg.max_smem_sm;
    int nw = env_int("QEB_NWZ", pick_warps<PL>(4, occ4 ? 5 : (stage ? 8 : 10)));
    if (occ4 && nw > 5) nw = 5;
    if (stage && nw > 8) nw = 8;
    if (inv) {
        if (occ4)
            k_vloc_zinv_f<PL, 4><<<grid, nw * 32, smem, st>>>(t, io);
        else
            k_vloc_zinv_f<PL, 1><<<grid, nw * 32, smem, st>>>(t, io);
    } else {
        if (stage && occ4)
            k_vloc_zfwd_f<PL, true, 4><<<grid, nw * 32, smem, st>>>(t, io);
        else if (stage)
            k_vloc_zfwd_f<PL, true, 1><<<grid, nw * 32, smem, st>>>(t, io);
        else
            k_vloc_zfwd_f<PL, false, 1><<<grid, nw * 32, smem, st>>>(t, io);
    }
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
template <class PL>
void FastLaunch<PL>::xspin(const VlocTables& t, const VlocLaunchCfg& cfg, int nbands, cudaStream_t st)
{
    if (env_int("QEB_XSPINX", 1)) {  // half-warp layout: one tile per spinor, 16 lines
        const int nw = env_int("QEB_NWS", pick_warps<PL>(4, 10));
        const dim3 gridx(nbands, (t.npos + 15) / 16);
        k_vloc_xspinx_f<PL><<<gridx, nw * 32, (size_t)t.nr1 * QEB_TL * 16 + PL::N * 20 + 16, st>>>(t);
        return;
    }
    const dim3 grid(nbands, t.ntx);
    // two tiles: one CTA per SM at the large grids, so the CTA carries all the warps the register file allows
    // (AUSURF112, 400 spinors: 10 warps 33.5 ms, 16 warps 31.0 ms)
    const size_t smem = 2 * cfg.smem_x + PL::N * 20 + 16;
    const int nw = env_int("QEB_NWS", 2 * smem > cfg.max_smem ? QEB_FAST_MAXTHREADS / 32 : pick_warps<PL>(4, 10));
    k_vloc_xspin_f<PL><<<grid, nw * 32, 2 * cfg.smem_x + PL::N * 20 + 16, st>>>(t);
}
template <class PL>
void FastLaunch<PL>::xacc(const VlocTables& t, const VlocIO& io, const VlocLaunchCfg& cfg, int nboxes, const double* w,
                          double* rho, cudaStream_t st)
{
    const int nw = env_int("QEB_NWX", pick_warps<PL>(4, 10));
    const size_t smem = cfg.smem_x + (size_t)t.nr1 * t.TX * 8 + PL::N * 24 + 32;
    // box chunks: enough CTAs for ~4 waves of the resident slots
    int per_sm = (int)(cfg.max_smem / smem);
    if (per_sm < 1) per_sm = 1;
    if (per_sm > 2048 / (nw * 32)) per_sm = 2048 / (nw * 32);
    int chunks = (int)((4L * cfg.num_sms * per_sm + t.ntx - 1) / t.ntx);
    if (chunks > nboxes) chunks = nboxes;
    if (chunks < 1) chunks = 1;
    const int bpc = (nboxes + chunks - 1) / chunks;
    const dim3 grid(t.ntx, (nboxes + bpc - 1) / bpc);
    k_vloc_xacc_f<PL><<<grid, nw * 32, smem, st>>>(t, io, w, nboxes, bpc, rho);
}
#endif

}  // namespace qeb
