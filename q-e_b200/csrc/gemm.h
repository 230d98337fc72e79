// Host launchers of the projector kernels (gemm.cu).
#pragma once
#include <cuda_runtime.h>
#include "common.cuh"
namespace qeb {
void calbec_plan(int n, int nkb, int m, bool cplx_mode, int num_sms, int* nsplit, int* gchunk);
size_t calbec_workspace_bytes(int nkb, int m, int nsplit, bool cplx_mode);
cudaError_t launch_calbec(const cplx* vkb, int npwx, const cplx* psi, int lda, int n, int nkb, int m, bool cplx_mode,
                          int gstart, int nsplit, int gchunk, void* part, void* becp, cudaStream_t st);
// noncollinear: m spinors = 2m columns of becp / ps (2b up, 2b+1 down); deeq_nc complex (nhm,nhm,nat,4)
cudaError_t launch_apply_deeq_nc(const void* becp, int nkb, int m, const int* row_atom, const int* ofs, const int* nh,
                                 const void* deeq_nc, int nhm, int nat, void* ps, int ldps, cudaStream_t st);
cudaError_t launch_apply_deeq(const void* becp, int nkb, int m, bool cplx_mode, const int* row_atom, const int* ofs,
                              const int* nh, const double* deeq, int nhm, void* ps, int ldps, cudaStream_t st);
cudaError_t launch_addvus(const cplx* vkb, int npwx, const void* ps, int ldps, int n, int nkb, int m, bool cplx_mode,
                          cplx* hpsi, int lda, cudaStream_t st);
cudaError_t launch_zero_im_g0(cplx* hpsi, int lda, int m, cudaStream_t st);
}  // namespace qeb
