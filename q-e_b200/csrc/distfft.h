// Slab ("R&G") distributed V_loc: stick <-> plane exchange between the z passes (own sticks, all z) and the
// y/x passes (all wave sticks, own z planes). Reference: FFTXlib/src/fft_scatter_2d.f90:38-266 (pack, mpi_alltoall,
// unpack through ismap), FFTXlib/src/fft_parallel_2d.f90:36-204 (cft_1z / fft_scatter / cft_2xy order).
#pragma once
#include <cuda_runtime.h>
#include <vector>
#include "common.cuh"
#include "nccl_glue.h"

namespace qeb {

struct DistDev {          // device tables of the exchange
    int nproc, me, nz_me, nst_tot, nst_me;
    const int* own;       // [nst_tot] owner rank of the plane-ordered wave stick sg
    const int* sidx;      // [nst_tot] position of sg in its owner's stick list
    const int* nsw;       // [nproc]   wave sticks per rank
    const int* prefix;    // [nproc]   wave sticks on lower ranks
};

struct DistHost {
    int nproc = 1, me = 0;
    std::vector<int> nr3p, i0r3p, nsw, prefix;
    int nst_tot = 0;
};

// R[rank p block][zl][box][s_p]  ->  G1p[box][zl][sg]      (after the forward exchange)
cudaError_t dist_unpack(const DistDev& d, const cplx* R, cplx* g1p, int nb, cudaStream_t st);
// G1p[box][zl][sg]  ->  R[rank p block][zl][box][s_p]      (before the backward exchange)
cudaError_t dist_pack(const DistDev& d, const cplx* g1p, cplx* R, int nb, cudaStream_t st);
// forward: S[z][box][s_me] (own sticks, all z) -> R ; backward: R -> S. One grouped send/recv pair per peer on `st`.
// Returns 0 or an NCCL error code (message through *msg).
int dist_exchange(const DistHost& h, NcclComm comm, bool forward, cplx* S, cplx* R, int nb, cudaStream_t st,
                  const char** msg);

// Peer-to-peer form of the same exchange (one process per GPU, buffers shared with cudaIpc*): every rank WRITES its
// blocks straight into the peers' R (forward) / S (backward) buffers with copy-engine transfers over NVLink, then
// raises a flag word in each peer's memory; the consumer's stream waits on its own flag words. No NCCL kernel and no
// SM is used for the transfer, and the copies of one batch overlap the FFT kernels of the other V_loc stream.
struct P2PState {
    bool ok = false;
    int nslots = 0, batch = 0;
    std::vector<cplx*> peer_S, peer_R;      // base of each rank's sbuf / rbuf (own entries: local pointers)
    std::vector<unsigned*> peer_flags;      // base of each rank's flag words [2 dirs][nslots][nproc]
    unsigned** peer_flags_dev = nullptr;    // device copy of peer_flags
    unsigned epoch[2][4] = {{0, 0, 0, 0}, {0, 0, 0, 0}};
};
// slot: V_loc stream index (each has its own S/R slice); batch_cap: boxes per slice (the same on every rank)
int dist_exchange_p2p(const DistHost& h, P2PState& p, bool forward, int slot, cplx* S, cplx* R, int nb, int nr3,
                      cudaStream_t st, const char** msg);

// signal + wait only (the fused forward leg: the z pass has already written the peers' R buffers)
int dist_p2p_sync(const DistHost& h, P2PState& p, bool forward, int slot, cudaStream_t st, const char** msg);

}  // namespace qeb
