// Device-side producers of the per-k-point state h_psi reads: g2_kin and init_us_2.
#pragma once
#include <cuda_runtime.h>
#include "common.cuh"
namespace qeb {
// g2kin(i) = |xk + g(:,igk(i))|^2 tpiba2 [+ qcutz (1 + erf((g2kin - ecfixed)/q2sigma))]   PW/src/g2_kin.f90:37-58
cudaError_t launch_g2_kin(int npw, const int* igk, const double* g, double xk1, double xk2, double xk3, double tpiba2,
                          double qcutz, double ecfixed, double q2sigma, double* g2kin, cudaStream_t st);
struct UsppTables {      // uspp / uspp_param / beta_mod module state (device pointers)
    int nsp, nhm, nbetam, lmaxkb, nqx;
    double dq;
    const int* nh;       // [nsp]
    const int* nhtol;    // [nhm*nsp]  l of beta ih           (Fortran nhtol(nhm,nsp))
    const int* nhtolm;   // [nhm*nsp]  combined lm, 1-based   (nhtolm(nhm,nsp))
    const int* indv;     // [nhm*nsp]  radial beta, 1-based   (indv(nhm,nsp))
    const double* tab_beta;  // [nqx*nbetam*nsp]             (tab_beta(nqx,nbetam,nsp))
};
// step 1 (upflib/init_us_2_acc.f90:92-108 + ylmr2.f90 + beta_mod.f90:130-173): per plane wave the (lmaxkb+1)^2 real
// spherical harmonics of q+G and the interpolated radial betas of every type: ylm[lm*npw+ig], vq[(nt*nbetam+nb)*npw+ig]
cudaError_t launch_us2_ylm_vq(int npw, const int* igk, const double* g, double q1, double q2, double q3, double tpiba,
                              const UsppTables& u, double* ylm, double* vq, cudaStream_t st);
// step 2 (:111-160): vkb(ig, jkb0(na)+ih) = ylm(ig,nhtolm) vq(ig,indv) (-i)^l eigts1 eigts2 eigts3 qphase(na), qphase = exp(-i 2pi q.tau)
cudaError_t launch_us2_vkb(int npw, int npwx, const int* igk, const int* mill, int nat, const int* ityp,
                           const int* jkb0, const cplx* qphase, int nr1, int nr2, int nr3, const cplx* eigts1,
                           const cplx* eigts2, const cplx* eigts3, const UsppTables& u, const double* ylm,
                           const double* vq, cplx* vkb, cudaStream_t st);
}  // namespace qeb
