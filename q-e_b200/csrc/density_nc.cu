// qeb_sum_band_rho_nc: density and magnetisation of noncollinear spinors, the `noncolin` branch of the band loop of
// sum_band (PW/src/sum_band.f90:555-575: two wave_g2r per band, get_rho_k on both components :650-671, get_rho_domag
// :699-729).
//
// Built on the scalar density path (qeb_sum_band_rho: scatter + z pass, y pass, x transform + accumulation, also on the
// slab-distributed descriptor), which accumulates sum_b w_b |f_b(r)|^2 of any set of columns f_b. With u, d the two
// components of a spinor and A = sum w|u|^2, B = sum w|d|^2, C = sum w|u + d|^2, D = sum w|u - i d|^2:
//     rho += A + B                      m_x += 2w Re(conj(u) d) = C - (A + B)
//     m_z += A - B                      m_y += 2w Im(conj(u) d) = D - (A + B)
// (the transform is linear, so u + d and u - i d are formed on the plane-wave coefficients). Without magnetisation
// (domag = 0) only rho(:,1) is touched and the 2m columns go through the scalar path in one call. With it the cost is
// four pruned transforms per band instead of the reference's two; a density kernel that pairs the two boxes of a spinor
// would halve that.
#include "context.h"

namespace {

// C(:, b) = u_b + d_b, D(:, b) = u_b - i d_b on rows < n; evc holds band b as columns 2b (up) and 2b + 1 (down) of lda rows
__global__ void k_nc_combine(const cplx* __restrict__ evc, int lda, int n, int m, cplx* __restrict__ C, cplx* __restrict__ D)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int b = blockIdx.y;
    if (i >= n || b >= m) return;
    const cplx u = evc[(size_t)(2 * b) * lda + i];
    const cplx d = evc[(size_t)(2 * b + 1) * lda + i];
    C[(size_t)b * lda + i] = mk(u.x + d.x, u.y + d.y);
    D[(size_t)b * lda + i] = mk(u.x + d.y, u.y - d.x);
}

// rho(nnr,4) += the increments assembled from t = (A, B, C, D), each nnr long
__global__ void k_nc_fold(double* __restrict__ rho, size_t nnr, const double* __restrict__ t)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nnr; i += (size_t)gridDim.x * blockDim.x) {
        const double a = t[i], b = t[nnr + i], cc = t[2 * nnr + i], dd = t[3 * nnr + i];
        const double s = a + b;
        rho[i] += s;
        rho[nnr + i] += cc - s;
        rho[2 * nnr + i] += dd - s;
        rho[3 * nnr + i] += a - b;
    }
}

__global__ void k_nc_weights(const double* __restrict__ w, int m, double* __restrict__ w2)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < m) w2[2 * b] = w2[2 * b + 1] = w[b];
}

}  // namespace

extern "C" int qeb_sum_band_rho_nc(qeb_ctx* c, int lda, int n, int m, const void* evc, const double* w, double* rho, int domag)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    int rc = check_call(c, lda, n, m, evc, rho, "qeb_sum_band_rho_nc");
    if (rc) return rc;
    if (!w) QEB_FAIL(c, 97, "qeb_sum_band_rho_nc: null weights");
    if (c->gamma) QEB_FAIL(c, 98, "qeb_sum_band_rho_nc: noncollinear spinors on a gamma_only grid (the reference excludes it)");
    if (m == 0) return 0;
    const size_t nnr = (size_t)c->nr1 * c->nr2 * (c->dist ? c->Tp.nr3 : c->nr3);
    const int ncomp = domag ? 4 : 1;
    const void* evc_d;
    if ((rc = stage_in(c, evc, (size_t)2 * lda * m * 16, c->psi_stage, &evc_d))) return rc;
    const void* w_d;
    if ((rc = stage_in(c, w, (size_t)m * 8, c->nc_w_d, &w_d))) return rc;
    // device layout of nc_rho_d: [A B C D | staged rho (host callers)]
    const bool rho_dev = is_device_ptr(rho);
    QEB_CUDA(c, c->nc_rho_d.ensure((4 + (rho_dev ? 0 : 4)) * nnr * 8));
    double* t = (double*)c->nc_rho_d.p;
    double* r = rho;
    if (!rho_dev) {
        r = t + 4 * nnr;
        QEB_CUDA(c, cudaMemcpyAsync(r, rho, ncomp * nnr * 8, cudaMemcpyHostToDevice, c->stream));
    }
    const bool was_async = c->async;
    c->async = 1;  // the pieces below run back to back on the context's stream; one synchronisation at the end
    QEB_CUDA(c, c->nc_w2_d.ensure((size_t)2 * m * 8));
    double* w2 = (double*)c->nc_w2_d.p;
    k_nc_weights<<<(m + 255) / 256, 256, 0, c->stream>>>((const double*)w_d, m, w2);
    if (!domag) {
        rc = qeb_sum_band_rho(c, lda, n, 2 * m, evc_d, w2, r);
    } else {
        QEB_CUDA(c, c->nc_comb_d.ensure((size_t)2 * lda * m * 16));
        cplx* C = (cplx*)c->nc_comb_d.p;
        cplx* D = C + (size_t)lda * m;
        dim3 g((n + 255) / 256, m);
        k_nc_combine<<<g, 256, 0, c->stream>>>((const cplx*)evc_d, lda, n, m, C, D);
        cudaError_t e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaMemsetAsync(t, 0, 4 * nnr * 8, c->stream);
        if (e != cudaSuccess) {
            c->async = was_async;
            QEB_CUDA(c, e);
        }
        // the up (down) components are the columns of a (2 lda, m) array starting at row 0 (lda)
        rc = qeb_sum_band_rho(c, 2 * lda, n, m, evc_d, (const double*)w_d, t);
        if (!rc) rc = qeb_sum_band_rho(c, 2 * lda, n, m, (const cplx*)evc_d + lda, (const double*)w_d, t + nnr);
        if (!rc) rc = qeb_sum_band_rho(c, lda, n, m, C, (const double*)w_d, t + 2 * nnr);
        if (!rc) rc = qeb_sum_band_rho(c, lda, n, m, D, (const double*)w_d, t + 3 * nnr);
        if (!rc) {
            int blocks = (int)((nnr + 255) / 256);
            if (blocks > c->num_sms * 8) blocks = c->num_sms * 8;
            k_nc_fold<<<blocks, 256, 0, c->stream>>>(r, nnr, t);
        }
    }
    c->async = was_async;
    if (rc) return rc;
    QEB_CUDA(c, cudaGetLastError());
    if (!rho_dev) {
        QEB_CUDA(c, cudaMemcpyAsync(rho, r, ncomp * nnr * 8, cudaMemcpyDeviceToHost, c->stream));
        QEB_CUDA(c, cudaStreamSynchronize(c->stream));
        return 0;
    }
    if (!is_device_ptr(evc) || !is_device_ptr(w)) QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    return finish(c);
}
