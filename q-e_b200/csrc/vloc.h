// Host launchers of the V_loc pass kernels (vloc.cu).
#pragma once
#include <cuda_runtime.h>
#include "vloc_kernels.cuh"
namespace qeb {
struct VlocLaunchCfg {
    size_t smem_z, smem_y, smem_x;
    int threads_z, threads_y, threads_x;
    // fused plane pass (vloc_plane.cuh): used when t.pl_v is set and smem_plane <= max_smem
    size_t smem_plane = 0, max_smem = 0;
    size_t max_smem_sm = 0;  // shared memory of one SM (occupancy planning of the z passes)
    int num_sms = 148;
};
cudaError_t vloc_init_attrs(size_t max_smem);
// one batch of nboxes boxes (bands for k, band pairs for Gamma)
cudaError_t vloc_launch_batch(const VlocTables& t, const VlocIO& io, const VlocLaunchCfg& cfg, int nboxes,
                              cudaStream_t st);
// noncollinear batch of nbands spinors (2*nbands boxes: components of band b are boxes 2b, 2b+1 = columns 2b, 2b+1 of
// io.psi / io.hpsi with leading dimension lda) through the 2x2 potential t.vnc: z | y | spinor x pass | y | z
cudaError_t vloc_launch_batch_nc(const VlocTables& t, const VlocIO& io, const VlocLaunchCfg& cfg, int nbands,
                                 cudaStream_t st);
// the pieces of a batch, used separately by the slab-distributed path (an exchange sits between them):
// z pass (inv: scatter + z inverse -> G1; !inv: G1 -> z forward + gather) and the three plane passes (y, fused x, y)
cudaError_t vloc_launch_zpass(bool inv, const VlocTables& t, const VlocIO& io, const VlocLaunchCfg& cfg, int nboxes,
                              cudaStream_t st);
cudaError_t vloc_launch_planes(const VlocTables& t, const VlocLaunchCfg& cfg, int nboxes, cudaStream_t st);
// vp[xpos*npos + y*nr3 + z] = v[x = rev_x[xpos] + nr1*(y + nr2*z)] * scale
cudaError_t vloc_prepare_v(const double* v, double* vp, const int* rev_x, int nr1, int nr2, int nr3, double scale,
                           cudaStream_t st);
// sum_band density accumulation for one batch: scatter + z inverse, y inverse, x inverse + rho += w |psi(r)|^2
// (w_dev: the m band weights wg/omega on the device, indexed by io.band0 + box; rho_dev: nr1*nr2*nr3, x fastest)
// planes_only: the y and x passes alone (slab-distributed path: the z pass and the exchange ran before, t = plane tables)
cudaError_t vloc_launch_density(const VlocTables& t, const VlocIO& io, const VlocLaunchCfg& cfg, int nboxes,
                                const double* w_dev, double* rho_dev, cudaStream_t st, bool planes_only = false);
// plane-kernel layout of the potential: vpl[(z*nr1 + xpos)*nr2 + ypos] = v[rev_x[xpos] + nr1*(rev_y[ypos] + nr2*z)] * scale
cudaError_t vloc_prepare_v_plane(const double* v, double* vpl, const int* rev_x, const int* rev_y, int nr1, int nr2,
                                 int nr3, double scale, cudaStream_t st);
// true when an instantiated plane kernel exists for these tables' (x,y) plans
bool vloc_plane_available(const VlocTables& t);
// true when vloc_launch_planes would run a fused plane kernel (not the y | x | y kernels) for these tables
bool vloc_uses_plane_kernel(const VlocTables& t, const VlocLaunchCfg& cfg);
}  // namespace qeb
