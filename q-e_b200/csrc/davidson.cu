#include "davidson.h"

namespace qeb {

__global__ void k_g_psi(int lda, int n, int m, cplx* __restrict__ psi, const double* __restrict__ e,
                        const double* __restrict__ h_diag, const double* __restrict__ s_diag)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int k = blockIdx.y;
    if (i >= n || k >= m) return;
    const double scala = 1.0;
    const double x = (h_diag[i] - e[k] * s_diag[i]) * scala;
    const double denm = 0.5 * (1.0 + x + sqrt(1.0 + (x - 1.0) * (x - 1.0))) / scala;
    cplx* p = psi + (size_t)k * lda + i;
    const cplx v = *p;
    *p = mk(v.x / denm, v.y / denm);
}

cudaError_t launch_g_psi(int lda, int n, int m, cplx* psi, const double* e, const double* h_diag, const double* s_diag,
                         cudaStream_t st)
{
    if (n <= 0 || m <= 0) return cudaSuccess;
    k_g_psi<<<dim3((n + 255) / 256, m), 256, 0, st>>>(lda, n, m, psi, e, h_diag, s_diag);
    return cudaGetLastError();
}

// npol = 2 (g_psi.f90:44-62): psi(lda, 2, m), h_diag / s_diag (lda, 2); column j of the (lda, 2m) view is component
// j & 1 of band j >> 1
__global__ void k_g_psi_nc(int lda, int n, int m, cplx* __restrict__ psi, const double* __restrict__ e,
                           const double* __restrict__ h_diag, const double* __restrict__ s_diag)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= n || j >= 2 * m) return;
    const int ipol = j & 1, k = j >> 1;
    const double x = h_diag[(size_t)ipol * lda + i] - e[k] * s_diag[(size_t)ipol * lda + i];
    const double denm = 0.5 * (1.0 + x + sqrt(1.0 + (x - 1.0) * (x - 1.0)));
    cplx* p = psi + (size_t)j * lda + i;
    const cplx v = *p;
    *p = mk(v.x / denm, v.y / denm);
}

cudaError_t launch_g_psi_nc(int lda, int n, int m, cplx* psi, const double* e, const double* h_diag, const double* s_diag,
                            cudaStream_t st)
{
    if (n <= 0 || m <= 0) return cudaSuccess;
    k_g_psi_nc<<<dim3((n + 255) / 256, 2 * m), 256, 0, st>>>(lda, n, m, psi, e, h_diag, s_diag);
    return cudaGetLastError();
}

__global__ void __launch_bounds__(256) k_col_dot(int n, const cplx* __restrict__ a, int lda, const cplx* __restrict__ b,
                                                 int ldb, int gamma, int gstart, double* __restrict__ out)
{
    __shared__ double red[256];
    const int k = blockIdx.x;
    const cplx* pa = a + (size_t)k * lda;
    const cplx* pb = b + (size_t)k * ldb;
    double s = 0.0;
    for (int i = threadIdx.x; i < n; i += 256) s += pa[i].x * pb[i].x + pa[i].y * pb[i].y;
    red[threadIdx.x] = s;
    __syncthreads();
    for (int w = 128; w > 0; w >>= 1) {
        if (threadIdx.x < w) red[threadIdx.x] += red[threadIdx.x + w];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        double r = red[0];
        if (gamma) {
            r = 2.0 * r;
            if (gstart == 2) r -= pa[0].x * pb[0].x;
        }
        out[k] = r;
    }
}

cudaError_t launch_col_dot(int n, int m, const cplx* a, int lda, const cplx* b, int ldb, int gamma, int gstart,
                           double* out, cudaStream_t st)
{
    if (m <= 0) return cudaSuccess;
    k_col_dot<<<m, 256, 0, st>>>(n, a, lda, b, ldb, gamma, gstart, out);
    return cudaGetLastError();
}

// rows n..lda-1 of m columns = 0 (h_psi.f90:141). A kernel, not cudaMemset2D: the pitched region of a 2-D memset that
// starts n elements into the first column ends n elements past the last one, i.e. outside an exactly sized array.
__global__ void k_zero_pad_rows(cplx* __restrict__ a, int lda, int n, int m)
{
    const int w = lda - n;
    const size_t tot = (size_t)w * m;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < tot; e += (size_t)gridDim.x * blockDim.x) {
        const size_t k = e / w;
        a[k * lda + n + (e - k * w)] = mk(0.0, 0.0);
    }
}

cudaError_t launch_zero_pad_rows(cplx* a, int lda, int n, int m, cudaStream_t st)
{
    if (m <= 0 || lda <= n) return cudaSuccess;
    const size_t tot = (size_t)(lda - n) * m;
    const int blocks = (int)((tot + 255) / 256 < 1184 ? (tot + 255) / 256 : 1184);
    k_zero_pad_rows<<<blocks, 256, 0, st>>>(a, lda, n, m);
    return cudaGetLastError();
}

}  // namespace qeb
