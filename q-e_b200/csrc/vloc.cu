// CUDA kernels of the V_loc path: thin wrappers that run the barrier-separated
// phases of vloc_kernels.cuh. See that file for the algorithm and the
// reference lines it replaces.
#include "vloc.h"
#include "vloc_fast_inst.cuh"
#include "vloc_plane_inst.cuh"
#include "vloc_plane2_inst.cuh"

namespace qeb {


__global__ void __launch_bounds__(256) k_vloc_zinv(const __grid_constant__ VlocTables t, const __grid_constant__ VlocIO io)
{
    QEB_SMEM;
    const int nph = zinv_nphase(t);
    for (int ph = 0; ph < nph; ++ph) {
        zinv_phase(ph, threadIdx.x, blockDim.x, t, io, blockIdx.x, blockIdx.y, smem);
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256) k_vloc_zfwd(const __grid_constant__ VlocTables t, const __grid_constant__ VlocIO io)
{
    QEB_SMEM;
    const int nph = zfwd_nphase(t);
    for (int ph = 0; ph < nph; ++ph) {
        zfwd_phase(ph, threadIdx.x, blockDim.x, t, io, blockIdx.x, blockIdx.y, smem);
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256) k_vloc_yinv(const __grid_constant__ VlocTables t)
{
    QEB_SMEM;
    const int nph = ypass_nphase(t);
    for (int ph = 0; ph < nph; ++ph) {
        yinv_phase(ph, threadIdx.x, blockDim.x, t, blockIdx.x, blockIdx.y, blockIdx.z, smem);
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256) k_vloc_yfwd(const __grid_constant__ VlocTables t)
{
    QEB_SMEM;
    const int nph = ypass_nphase(t);
    for (int ph = 0; ph < nph; ++ph) {
        yfwd_phase(ph, threadIdx.x, blockDim.x, t, blockIdx.x, blockIdx.y, blockIdx.z, smem);
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256) k_vloc_xfus(const __grid_constant__ VlocTables t)
{
    QEB_SMEM;
    const int nph = xfus_nphase(t);
    for (int ph = 0; ph < nph; ++ph) {
        xfus_phase(ph, threadIdx.x, blockDim.x, t, blockIdx.x, blockIdx.y, smem);
        __syncthreads();
    }
}



// spinor x pass (noncollinear 2x2 potential): block = (line tile, band), two shared-memory tiles
__global__ void __launch_bounds__(256) k_vloc_xspin(const __grid_constant__ VlocTables t)
{
    QEB_SMEM;
    const int nph = xspin_nphase(t);
    for (int ph = 0; ph < nph; ++ph) {
        xspin_phase(ph, threadIdx.x, blockDim.x, t, blockIdx.x, blockIdx.y, smem);
        __syncthreads();
    }
}

// density accumulation: x inverse of every box of the batch + |psi(r)|^2 weights, one CTA per line tile
__global__ void __launch_bounds__(256) k_vloc_xacc(const __grid_constant__ VlocTables t, const __grid_constant__ VlocIO io,
                                                   const double* __restrict__ w, int nboxes, double* __restrict__ rho)
{
    QEB_SMEM;
    double* racc = reinterpret_cast<double*>(smem + (size_t)t.nr1 * t.TX);
    int* irev = reinterpret_cast<int*>(racc + (size_t)t.nr1 * t.TX);
    for (int i = threadIdx.x; i < t.nr1; i += blockDim.x) irev[t.px.rev[i]] = i;
    xacc_zero_phase(threadIdx.x, blockDim.x, t, racc);
    __syncthreads();
    const int np = t.px.npass;
    for (int box = 0; box < nboxes; ++box)
        for (int ph = 0; ph < np; ++ph) {
            xacc_phase(ph, threadIdx.x, blockDim.x, t, io, w, blockIdx.x, box, smem, racc);
            __syncthreads();
        }
    xacc_flush_phase(threadIdx.x, blockDim.x, t, blockIdx.x, irev, racc, rho);
}

// ---- fast variants (vloc_fast_inst.cuh, instantiated in vloc_fast_<N>.cu) ----
#define QEB_FOR_PLANS(X) X(SP15) X(SP25) X(SP27) X(SP64) X(SP125) X(SP135) X(SP144) X(SP200) X(SP200b) X(SP200c) X(SP200d)
#define X(PL) extern template struct FastLaunch<PL>;
QEB_FOR_PLANS(X)
#undef X

static bool launch_z(bool inv, const VlocTables& t, const VlocIO& io, const VlocLaunchCfg& cfg, int nboxes, cudaStream_t st)
{
    if (t.TZ > QEB_TL || env_int("QEB_GENERIC", 0)) return false;
#define X(PL)                                        \
    if (PL::matches(t.pz)) {                         \
        FastLaunch<PL>::z(inv, t, io, cfg, nboxes, st); \
        return true;                                 \
    }
    QEB_FOR_PLANS(X)
#undef X
    return false;
}
static bool launch_y(bool inv, const VlocTables& t, const VlocLaunchCfg& cfg, int nboxes, cudaStream_t st)
{
    if (t.TY > QEB_TL || env_int("QEB_GENERIC", 0)) return false;
#define X(PL)                                    \
    if (PL::matches(t.py)) {                     \
        FastLaunch<PL>::y(inv, t, cfg, nboxes, st); \
        return true;                             \
    }
    QEB_FOR_PLANS(X)
#undef X
    return false;
}
static bool launch_x(const VlocTables& t, const VlocLaunchCfg& cfg, int nboxes, cudaStream_t st)
{
    if (t.TX > QEB_TL || env_int("QEB_GENERIC", 0)) return false;
#define X(PL)                               \
    if (PL::matches(t.px)) {                \
        FastLaunch<PL>::x(t, cfg, nboxes, st); \
        return true;                        \
    }
    QEB_FOR_PLANS(X)
#undef X
    return false;
}

// ---- fused plane pass (vloc_plane_inst.cuh, instantiated in vloc_plane_<n1>_<n2>.cu) ----
#define QEB_FOR_PLANES(X) \
    X(SP125, SP64, 512) X(SP25, SP25, 512) X(SP27, SP27, 512) X(SP15, SP15, 512) X(SP64, SP64, 512) \
    X(SP125b, SP64, 1024) X(SP125b, SP64b, 1024)
#define X(PX, PY, MT) extern template struct PlaneLaunch<PX, PY, MT>;
QEB_FOR_PLANES(X)
#undef X

// second form of the plane kernel (vloc_plane2.cuh: compile-time phases, persistent over the (box, z) list, bulk
// asynchronous stick rows); preferred when instantiated for the plans (QEB_PLANE2=0: the first form)
#define QEB_FOR_PLANE2S(X) \
    X(SP125, SP64, 512) X(SP125b, SP64, 800) X(SP25, SP25, 512) X(SP27, SP27, 512) X(SP15, SP15, 512) X(SP64, SP64, 512)
#define X(PX, PY, NT) extern template struct Plane2Launch<PX, PY, NT>;
QEB_FOR_PLANE2S(X)
#undef X

static bool launch_plane2(const VlocTables& t, const VlocLaunchCfg& cfg, int nboxes, cudaStream_t st)
{
    if (!env_int("QEB_PLANE2", 1)) return false;
#define X(PX, PY, NT) \
    if (Plane2Launch<PX, PY, NT>::run(t, nboxes, cfg.num_sms, cfg.max_smem, st)) return true;
    QEB_FOR_PLANE2S(X)
#undef X
    return false;
}

bool vloc_plane_available(const VlocTables& t)
{
#define X(PX, PY, MT) \
    if (PX::matches(t.px) && PY::matches(t.py)) return true;
    QEB_FOR_PLANES(X)
#undef X
    return false;
}

bool vloc_uses_plane_kernel(const VlocTables& t, const VlocLaunchCfg& cfg)
{
    return t.pl_v && cfg.smem_plane != 0 && cfg.smem_plane <= cfg.max_smem && !env_int("QEB_GENERIC", 0) && env_int("QEB_PLANE", 1) &&
           vloc_plane_available(t);
}

static bool launch_plane(const VlocTables& t, const VlocLaunchCfg& cfg, int nboxes, cudaStream_t st)
{
    if (!t.pl_v || cfg.smem_plane == 0 || cfg.smem_plane > cfg.max_smem || env_int("QEB_GENERIC", 0) ||
        !env_int("QEB_PLANE", 1))
        return false;
    if (launch_plane2(t, cfg, nboxes, st)) return true;
#define X(PX, PY, MT) \
    if (PlaneLaunch<PX, PY, MT>::run(t, nboxes, cfg.num_sms, cfg.max_smem, st)) return true;
    QEB_FOR_PLANES(X)
#undef X
    return false;
}

__global__ void k_prepare_v_plane(const double* __restrict__ v, double* __restrict__ vpl, const int* __restrict__ rev_x,
                                  const int* __restrict__ rev_y, int nr1, int nr2, int nr3, double scale)
{
    // one block per (xpos chunk, z): reads are a gather inside one (x,y) plane (L1/L2 resident), writes coalesced
    const int z = blockIdx.y;
    const size_t plane = (size_t)nr1 * nr2;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < (int)plane; i += gridDim.x * blockDim.x) {
        const int xp = i / nr2, yp = i - xp * nr2;
        vpl[(size_t)z * plane + i] = v[rev_x[xp] + (size_t)nr1 * (rev_y[yp] + (size_t)nr2 * z)] * scale;
    }
}

__global__ void k_prepare_v(const double* __restrict__ v, double* __restrict__ vp, const int* __restrict__ rev_x,
                            int nr1, int nr2, int nr3, double scale)
{
    // one block per (y, z-chunk); stage a [32 z][nr1 x] tile through shared memory so that both the
    // x-fastest reads of v and the z-fastest writes of vp are coalesced
    extern __shared__ double tile[];  // [32][nr1+1]
    const int y = blockIdx.y, z0 = blockIdx.x * 32;
    const int ld = nr1 + 1;
    for (int i = threadIdx.x; i < 32 * nr1; i += blockDim.x) {
        const int zl = i / nr1, x = i - zl * nr1;
        if (z0 + zl < nr3) tile[zl * ld + x] = v[x + (size_t)nr1 * (y + (size_t)nr2 * (z0 + zl))];
    }
    __syncthreads();
    const size_t npos = (size_t)nr2 * nr3;
    for (int i = threadIdx.x; i < 32 * nr1; i += blockDim.x) {
        const int xp = i / 32, zl = i - xp * 32;
        if (z0 + zl < nr3) vp[(size_t)xp * npos + (size_t)y * nr3 + z0 + zl] = tile[zl * ld + rev_x[xp]] * scale;
    }
}

cudaError_t vloc_init_attrs(size_t max_smem)
{
    cudaError_t e;
    const int s = (int)max_smem;
    if ((e = cudaFuncSetAttribute(k_vloc_zinv, cudaFuncAttributeMaxDynamicSharedMemorySize, s))) return e;
    if ((e = cudaFuncSetAttribute(k_vloc_zfwd, cudaFuncAttributeMaxDynamicSharedMemorySize, s))) return e;
    if ((e = cudaFuncSetAttribute(k_vloc_yinv, cudaFuncAttributeMaxDynamicSharedMemorySize, s))) return e;
    if ((e = cudaFuncSetAttribute(k_vloc_yfwd, cudaFuncAttributeMaxDynamicSharedMemorySize, s))) return e;
    if ((e = cudaFuncSetAttribute(k_vloc_xfus, cudaFuncAttributeMaxDynamicSharedMemorySize, s))) return e;
    if ((e = cudaFuncSetAttribute(k_vloc_xacc, cudaFuncAttributeMaxDynamicSharedMemorySize, s))) return e;
    if ((e = cudaFuncSetAttribute(k_vloc_xspin, cudaFuncAttributeMaxDynamicSharedMemorySize, s))) return e;
#define X(PL) FastLaunch<PL>::attrs(s);
    QEB_FOR_PLANS(X)
#undef X
#define X(PX, PY, MT) \
    if ((e = PlaneLaunch<PX, PY, MT>::attrs(s))) return e;
    QEB_FOR_PLANES(X)
#undef X
#define X(PX, PY, NT) \
    if ((e = Plane2Launch<PX, PY, NT>::attrs(s))) return e;
    QEB_FOR_PLANE2S(X)
#undef X
    return cudaGetLastError();
}

cudaError_t vloc_launch_zpass(bool inv, const VlocTables& t, const VlocIO& io, const VlocLaunchCfg& cfg, int nboxes,
                              cudaStream_t st)
{
    if (nboxes <= 0 || t.ngrp <= 0) return cudaSuccess;
    if (!launch_z(inv, t, io, cfg, nboxes, st)) {
        if (inv)
            k_vloc_zinv<<<dim3(t.ngrp, nboxes), cfg.threads_z, cfg.smem_z, st>>>(t, io);
        else
            k_vloc_zfwd<<<dim3(t.ngrp, nboxes), cfg.threads_z, cfg.smem_z, st>>>(t, io);
    }
    return cudaGetLastError();
}

cudaError_t vloc_launch_planes(const VlocTables& t, const VlocLaunchCfg& cfg, int nboxes, cudaStream_t st)
{
    if (nboxes <= 0 || t.nxw <= 0 || t.npos <= 0) return cudaSuccess;
    if (launch_plane(t, cfg, nboxes, st)) return cudaGetLastError();
    if (!launch_y(true, t, cfg, nboxes, st))
        k_vloc_yinv<<<dim3(t.nty, t.nxw, nboxes), cfg.threads_y, cfg.smem_y, st>>>(t);
    if (!launch_x(t, cfg, nboxes, st))
        k_vloc_xfus<<<dim3(t.ntx, nboxes), cfg.threads_x, cfg.smem_x, st>>>(t);
    if (!launch_y(false, t, cfg, nboxes, st))
        k_vloc_yfwd<<<dim3(t.nty, t.nxw, nboxes), cfg.threads_y, cfg.smem_y, st>>>(t);
    return cudaGetLastError();
}

cudaError_t vloc_launch_batch(const VlocTables& t, const VlocIO& io, const VlocLaunchCfg& cfg, int nboxes,
                              cudaStream_t st)
{
    if (nboxes <= 0) return cudaSuccess;
    cudaError_t e;
    if ((e = vloc_launch_zpass(true, t, io, cfg, nboxes, st)) != cudaSuccess) return e;
    if ((e = vloc_launch_planes(t, cfg, nboxes, st)) != cudaSuccess) return e;
    return vloc_launch_zpass(false, t, io, cfg, nboxes, st);
}

cudaError_t vloc_launch_batch_nc(const VlocTables& t, const VlocIO& io, const VlocLaunchCfg& cfg, int nbands,
                                 cudaStream_t st)
{
    if (nbands <= 0) return cudaSuccess;
    if (!t.vnc || 2 * cfg.smem_x > cfg.max_smem) return cudaErrorInvalidConfiguration;
    const int nboxes = 2 * nbands;
    cudaError_t e;
    if ((e = vloc_launch_zpass(true, t, io, cfg, nboxes, st)) != cudaSuccess) return e;
    if (t.nxw > 0 && t.npos > 0) {
        if (!launch_y(true, t, cfg, nboxes, st))
            k_vloc_yinv<<<dim3(t.nty, t.nxw, nboxes), cfg.threads_y, cfg.smem_y, st>>>(t);
        // static-plan kernel for the multi-pass plans (a single-pass plan would hold 2*N complex values per thread)
        bool done = false;
        if (t.TX <= QEB_TL && !env_int("QEB_GENERIC", 0) && 2 * cfg.smem_x + (size_t)t.nr1 * 20 + 16 <= cfg.max_smem) {
#define X(PL)                                            \
    if (!done && PL::NP > 1 && PL::matches(t.px)) {      \
        FastLaunch<PL>::xspin(t, cfg, nbands, st);       \
        done = true;                                     \
    }
            QEB_FOR_PLANS(X)
#undef X
        }
        if (!done) k_vloc_xspin<<<dim3(t.ntx, nbands), cfg.threads_x, 2 * cfg.smem_x, st>>>(t);
        if (!launch_y(false, t, cfg, nboxes, st))
            k_vloc_yfwd<<<dim3(t.nty, t.nxw, nboxes), cfg.threads_y, cfg.smem_y, st>>>(t);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
    }
    return vloc_launch_zpass(false, t, io, cfg, nboxes, st);
}

cudaError_t vloc_prepare_v(const double* v, double* vp, const int* rev_x, int nr1, int nr2, int nr3, double scale,
                           cudaStream_t st)
{
    dim3 grid((nr3 + 31) / 32, nr2);
    k_prepare_v<<<grid, 256, (size_t)32 * (nr1 + 1) * 8, st>>>(v, vp, rev_x, nr1, nr2, nr3, scale);
    return cudaGetLastError();
}

cudaError_t vloc_launch_density(const VlocTables& t, const VlocIO& io, const VlocLaunchCfg& cfg, int nboxes,
                                const double* w_dev, double* rho_dev, cudaStream_t st, bool planes_only)
{
    if (nboxes <= 0 || t.nxw <= 0) return cudaSuccess;
    cudaError_t e;
    if (!planes_only && (e = vloc_launch_zpass(true, t, io, cfg, nboxes, st)) != cudaSuccess) return e;
    if (!launch_y(true, t, cfg, nboxes, st))
        k_vloc_yinv<<<dim3(t.nty, t.nxw, nboxes), cfg.threads_y, cfg.smem_y, st>>>(t);
    const size_t smem = cfg.smem_x + (size_t)t.nr1 * t.TX * 8 + (size_t)t.nr1 * 4 + 16;
    if (smem + (size_t)t.nr1 * 20 > cfg.max_smem) return cudaErrorInvalidConfiguration;
    if (t.TX <= QEB_TL && !env_int("QEB_GENERIC", 0)) {
#define X(PL)                                                          \
    if (PL::matches(t.px)) {                                           \
        FastLaunch<PL>::xacc(t, io, cfg, nboxes, w_dev, rho_dev, st);  \
        return cudaGetLastError();                                     \
    }
        QEB_FOR_PLANS(X)
#undef X
    }
    k_vloc_xacc<<<t.ntx, cfg.threads_x, smem, st>>>(t, io, w_dev, nboxes, rho_dev);
    return cudaGetLastError();
}

cudaError_t vloc_prepare_v_plane(const double* v, double* vpl, const int* rev_x, const int* rev_y, int nr1, int nr2,
                                 int nr3, double scale, cudaStream_t st)
{
    const int per = (nr1 * nr2 + 255) / 256;
    dim3 grid(per < 64 ? per : 64, nr3);
    k_prepare_v_plane<<<grid, 256, 0, st>>>(v, vpl, rev_x, rev_y, nr1, nr2, nr3, scale);
    return cudaGetLastError();
}

}  // namespace qeb
