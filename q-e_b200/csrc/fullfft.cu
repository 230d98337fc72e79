// Full-box 3-D FFT kernels (qeb_fft): wrappers around fullfft_kernels.cuh.
#include "fullfft.h"
#include "fullfft_kernels.cuh"

namespace qeb {

#define QEB_SMEM extern __shared__ __align__(16) unsigned char smem_raw[]; cplx* smem = reinterpret_cast<cplx*>(smem_raw)

__global__ void __launch_bounds__(256) k_full_z(const __grid_constant__ FullFftParams p)
{
    QEB_SMEM;
    for (int ph = 0; ph < p.pz.npass; ++ph) {
        fullz_phase(ph, threadIdx.x, blockDim.x, p, blockIdx.x, blockIdx.y, smem);
        __syncthreads();
    }
}
__global__ void __launch_bounds__(256) k_full_y(const __grid_constant__ FullFftParams p)
{
    QEB_SMEM;
    for (int ph = 0; ph < p.py.npass; ++ph) {
        fully_phase(ph, threadIdx.x, blockDim.x, p, blockIdx.x, blockIdx.y, blockIdx.z, smem);
        __syncthreads();
    }
}
__global__ void __launch_bounds__(256) k_full_x(const __grid_constant__ FullFftParams p)
{
    QEB_SMEM;
    const int nph = fullx_nphase(p);
    for (int ph = 0; ph < nph; ++ph) {
        fullx_phase(ph, threadIdx.x, blockDim.x, p, blockIdx.x, blockIdx.y, smem);
        __syncthreads();
    }
}

static int tile_for(int n, size_t cap, int pad)
{
    int t = 32;
    while (t > 1 && (size_t)n * (t + pad) * 16 > cap) t -= 2;
    return t < 1 ? 1 : t;
}

cudaError_t fullfft_launch(const FftPlan& px, const FftPlan& py, const FftPlan& pz, cplx* f, int howmany, int sign,
                           size_t max_smem, cudaStream_t st)
{
    static bool attr = false;
    cudaError_t e;
    if (!attr) {
        if ((e = cudaFuncSetAttribute(k_full_z, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)max_smem))) return e;
        if ((e = cudaFuncSetAttribute(k_full_y, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)max_smem))) return e;
        if ((e = cudaFuncSetAttribute(k_full_x, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)max_smem))) return e;
        attr = true;
    }
    FullFftParams p;
    p.nr1 = px.n; p.nr2 = py.n; p.nr3 = pz.n;
    p.px = px; p.py = py; p.pz = pz;
    p.f = f;
    p.sgn = sign > 0 ? 1.0 : -1.0;
    p.scale = sign > 0 ? 1.0 : 1.0 / ((double)p.nr1 * p.nr2 * p.nr3);
    const size_t cap = 64 * 1024;
    const int ncol = p.nr1 * p.nr2, nrows = p.nr2 * p.nr3;
    p.TZ = tile_for(p.nr3, cap, 0); if (p.TZ > ncol) p.TZ = ncol;
    p.TYx = tile_for(p.nr2, cap, 0); if (p.TYx > p.nr1) p.TYx = p.nr1;
    p.TXr = tile_for(p.nr1, cap, 1); if (p.TXr > nrows) p.TXr = nrows;
    const int ntz = (ncol + p.TZ - 1) / p.TZ, nty = (p.nr1 + p.TYx - 1) / p.TYx, ntx = (nrows + p.TXr - 1) / p.TXr;
    p.dTZ = make_fastdiv(p.TZ); p.dTZl = make_fastdiv(ncol - (ntz - 1) * p.TZ);
    p.dTY = make_fastdiv(p.TYx); p.dTYl = make_fastdiv(p.nr1 - (nty - 1) * p.TYx);
    p.dTX = make_fastdiv(p.TXr); p.dTXl = make_fastdiv(nrows - (ntx - 1) * p.TXr);
    const size_t sz = (size_t)p.nr3 * p.TZ * 16, sy = (size_t)p.nr2 * p.TYx * 16, sx = (size_t)p.nr1 * (p.TXr + 1) * 16;
    if (sz > max_smem || sy > max_smem || sx > max_smem) return cudaErrorInvalidValue;
    k_full_z<<<dim3(ntz, howmany), 256, sz, st>>>(p);
    k_full_y<<<dim3(nty, p.nr3, howmany), 256, sy, st>>>(p);
    k_full_x<<<dim3(ntx, howmany), 256, sx, st>>>(p);
    return cudaGetLastError();
}

}  // namespace qeb
