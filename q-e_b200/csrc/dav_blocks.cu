// libqeb200, translation unit 6 of 6 (with solver.cu): the building blocks of the block-Davidson solver on device arrays -
// g_psi, column norms, the reduced-matrix and basis-update GEMMs on the DMMA projector kernels.
#include "context.h"

extern "C" {

// ---- block-Davidson building blocks ---------------------------
int qeb_set_g_psi_diag(qeb_ctx* c, int n, const double* h_diag, const double* s_diag)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (n < 1 || !h_diag || !s_diag) QEB_FAIL(c, 90, "qeb_set_g_psi_diag: bad arguments");
    QEB_CUDA(c, c->gpsi_d.ensure((size_t)n * 16));
    QEB_CUDA(c, cudaMemcpyAsync(c->gpsi_d.p, h_diag, (size_t)n * 8,
                                is_device_ptr(h_diag) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
    QEB_CUDA(c, cudaMemcpyAsync((double*)c->gpsi_d.p + n, s_diag, (size_t)n * 8,
                                is_device_ptr(s_diag) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
    QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    c->gpsi_n = n;
    return 0;
}

int qeb_g_psi(qeb_ctx* c, int lda, int n, int m, int npol, void* psi, const double* e)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (npol != 1 && npol != 2) QEB_FAIL(c, 91, "qeb_g_psi: npol must be 1 or 2");
    if (n < 1 || lda < n || m < 0 || !psi || !e) QEB_FAIL(c, 91, "qeb_g_psi: bad arguments");
    if (c->gpsi_n < n) QEB_FAIL(c, 92, "qeb_g_psi: qeb_set_g_psi_diag has not been called for this n");
    // npol = 2: the diagonals are h_diag(npwx, npol), s_diag(npwx, npol) with npwx = lda, set with n = npwx*npol
    if (npol == 2 && c->gpsi_n != 2 * lda)
        QEB_FAIL(c, 92, "qeb_g_psi: npol = 2 needs qeb_set_g_psi_diag with n = 2*lda (h_diag(npwx,npol), s_diag(npwx,npol))");
    if (!is_device_ptr(psi)) QEB_FAIL(c, 93, "qeb_g_psi: psi must be a device pointer (the Davidson work arrays live on the device)");
    if (m == 0) return 0;
    const void* ed;
    int rc = stage_in(c, e, (size_t)m * 8, c->dav_small_d, &ed);
    if (rc) return rc;
    if (npol == 2)
        QEB_CUDA(c, launch_g_psi_nc(lda, n, m, (cplx*)psi, (const double*)ed, (const double*)c->gpsi_d.p,
                                    (const double*)c->gpsi_d.p + c->gpsi_n, c->stream));
    else
        QEB_CUDA(c, launch_g_psi(lda, n, m, (cplx*)psi, (const double*)ed, (const double*)c->gpsi_d.p,
                                 (const double*)c->gpsi_d.p + c->gpsi_n, c->stream));
    if (!is_device_ptr(e)) QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    return finish(c);
}

int qeb_col_dot(qeb_ctx* c, int n, int m, const void* a, int lda, const void* b, int ldb, double* out)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (n < 1 || m < 0 || lda < n || ldb < n || !a || !b || !out) QEB_FAIL(c, 94, "qeb_col_dot: bad arguments");
    if (!is_device_ptr(a) || !is_device_ptr(b)) QEB_FAIL(c, 93, "qeb_col_dot: a and b must be device pointers");
    if (m == 0) return 0;
    const bool od = is_device_ptr(out);
    double* o = out;
    if (!od) {
        QEB_CUDA(c, c->dav_small_d.ensure((size_t)m * 8));
        o = (double*)c->dav_small_d.p;
    }
    QEB_CUDA(c, launch_col_dot(n, m, (const cplx*)a, lda, (const cplx*)b, ldb, c->gamma ? 1 : 0, c->gstart, o, c->stream));
    if (!od) {
        QEB_CUDA(c, cudaMemcpyAsync(out, o, (size_t)m * 8, cudaMemcpyDeviceToHost, c->stream));
        QEB_CUDA(c, cudaStreamSynchronize(c->stream));
        return 0;
    }
    return finish(c);
}

// C(ka,mb) = A(1:n,1:ka)^H B(1:n,1:mb): the reduced matrices hc/sc of cegterg (ZGEMM 'C','N', cegterg.f90:209-229,
// 478-499) and hr/sr of regterg (2*DGEMM('T','N') - DGER at G=0, regterg.f90:203-221, 402-429) on the calbec kernel.
int qeb_dav_gemm_cn(qeb_ctx* c, int n, const void* A, int lda, int ka, const void* B, int ldb, int mb, void* C, int ldc)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (!c->have_grid) QEB_FAIL(c, 95, "qeb_dav_gemm_cn: qeb_set_grid has not been called (gamma_only / gstart are context state)");
    if (n < 1 || lda < n || ldb < n || ka < 0 || mb < 0 || ldc < ka || !A || !B || !C) QEB_FAIL(c, 95, "qeb_dav_gemm_cn: bad arguments");
    if (!is_device_ptr(A) || !is_device_ptr(B)) QEB_FAIL(c, 93, "qeb_dav_gemm_cn: A and B must be device pointers");
    if (ka == 0 || mb == 0) return 0;
    const size_t el = c->gamma ? 8 : 16;
    int nsplit, gchunk;
    calbec_plan(n, ka, mb, !c->gamma, c->num_sms, &nsplit, &gchunk);
    QEB_CUDA(c, c->part_d.ensure(calbec_workspace_bytes(ka, mb, nsplit, !c->gamma)));
    QEB_CUDA(c, c->misc_stage.ensure((size_t)ka * mb * el));
    QEB_CUDA(c, launch_calbec((const cplx*)A, lda, (const cplx*)B, ldb, n, ka, mb, !c->gamma, c->gstart, nsplit, gchunk,
                              c->part_d.p, c->misc_stage.p, c->stream));
    QEB_CUDA(c, cudaMemcpy2DAsync(C, (size_t)ldc * el, c->misc_stage.p, (size_t)ka * el, (size_t)ka * el, mb,
                                  is_device_ptr(C) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, c->stream));
    if (!is_device_ptr(C)) {
        QEB_CUDA(c, cudaStreamSynchronize(c->stream));
        return 0;
    }
    return finish(c);
}

// C(1:n,1:m) (+)= A(1:n,1:k) B(1:k,1:m): the basis updates of cegterg (ZGEMM 'N','N', cegterg.f90:353-395, 588-632)
// and regterg (DGEMM on the 2*npwx real view, B real) on the add_vuspsi kernel. accumulate == 0: C is overwritten
// (rows n+1..ldc of C are left untouched, as ZGEMM does).
int qeb_dav_gemm_nn(qeb_ctx* c, int n, const void* A, int lda, int k, const void* B, int ldb, int m, void* C, int ldc,
                    int accumulate)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (!c->have_grid) QEB_FAIL(c, 96, "qeb_dav_gemm_nn: qeb_set_grid has not been called");
    if (n < 1 || lda < n || ldc < n || k < 0 || m < 0 || ldb < k || !A || !B || !C) QEB_FAIL(c, 96, "qeb_dav_gemm_nn: bad arguments");
    if (!is_device_ptr(A) || !is_device_ptr(C)) QEB_FAIL(c, 93, "qeb_dav_gemm_nn: A and C must be device pointers");
    if (m == 0) return 0;
    if (!accumulate) QEB_CUDA(c, cudaMemset2DAsync(C, (size_t)ldc * 16, 0, (size_t)n * 16, m, c->stream));
    if (k == 0) return finish(c);
    const size_t el = c->gamma ? 8 : 16;
    const int ldps = (k + 1) & ~1;
    QEB_CUDA(c, c->misc_stage.ensure((size_t)ldps * m * el));
    QEB_CUDA(c, cudaMemsetAsync(c->misc_stage.p, 0, (size_t)ldps * m * el, c->stream));
    QEB_CUDA(c, cudaMemcpy2DAsync(c->misc_stage.p, (size_t)ldps * el, B, (size_t)ldb * el, (size_t)k * el, m,
                                  is_device_ptr(B) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
    QEB_CUDA(c, launch_addvus((const cplx*)A, lda, c->misc_stage.p, ldps, n, k, m, !c->gamma, (cplx*)C, ldc, c->stream));
    if (!is_device_ptr(B)) QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    return finish(c);
}

}  // extern "C"
