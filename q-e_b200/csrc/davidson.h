// Element-wise / reduction kernels of the block-Davidson inner loop; the tall-skinny GEMMs of
// cegterg/regterg run on the projector DMMA kernels (gemm.h).
#pragma once
#include <cuda_runtime.h>
#include "common.cuh"
namespace qeb {
// g_psi (PW/src/g_psi.f90:10-70, TEST_NEW_PRECONDITIONING branch): psi(i,k) /= denm(h_diag(i) - e(k) s_diag(i))
cudaError_t launch_g_psi(int lda, int n, int m, cplx* psi, const double* e, const double* h_diag, const double* s_diag,
                         cudaStream_t st);
// the same for spinors: psi(lda, 2, m), h_diag / s_diag (lda, 2) (g_psi.f90:44-62 with npol = 2)
cudaError_t launch_g_psi_nc(int lda, int n, int m, cplx* psi, const double* e, const double* h_diag, const double* s_diag,
                            cudaStream_t st);
// out(k) = sum_i Re(conj(a(i,k)) b(i,k)) over i < n  (MYDDOT(2*npw, a_k, b_k), cegterg.f90:421); Gamma:
// 2*sum - [gstart==2] Re a(1,k) Re b(1,k) (regterg.f90:340-345). Fixed reduction order: deterministic.
cudaError_t launch_col_dot(int n, int m, const cplx* a, int lda, const cplx* b, int ldb, int gamma, int gstart,
                           double* out, cudaStream_t st);
// a(n+1:lda, 1:m) = 0: the pad rows h_psi zeroes (PW/src/h_psi.f90:141)
cudaError_t launch_zero_pad_rows(cplx* a, int lda, int n, int m, cudaStream_t st);
}  // namespace qeb
