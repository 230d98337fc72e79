#include "distfft.h"

namespace qeb {

template <bool UNPACK>
__global__ void __launch_bounds__(256) k_dist_repack(const DistDev d, const cplx* __restrict__ src, cplx* __restrict__ dst,
                                                      int nb)
{
    const int sg = blockIdx.x * blockDim.x + threadIdx.x;
    if (sg >= d.nst_tot) return;
    const int zl = blockIdx.y, box = blockIdx.z;
    const int p = d.own[sg];
    const size_t r = (size_t)nb * d.nz_me * d.prefix[p] + ((size_t)zl * nb + box) * d.nsw[p] + d.sidx[sg];
    const size_t g = ((size_t)box * d.nz_me + zl) * d.nst_tot + sg;
    if (UNPACK)
        dst[g] = src[r];
    else
        dst[r] = src[g];
}

cudaError_t dist_unpack(const DistDev& d, const cplx* R, cplx* g1p, int nb, cudaStream_t st)
{
    if (d.nst_tot <= 0 || d.nz_me <= 0 || nb <= 0) return cudaSuccess;
    dim3 grid((d.nst_tot + 255) / 256, d.nz_me, nb);
    k_dist_repack<true><<<grid, 256, 0, st>>>(d, R, g1p, nb);
    return cudaGetLastError();
}

cudaError_t dist_pack(const DistDev& d, const cplx* g1p, cplx* R, int nb, cudaStream_t st)
{
    if (d.nst_tot <= 0 || d.nz_me <= 0 || nb <= 0) return cudaSuccess;
    dim3 grid((d.nst_tot + 255) / 256, d.nz_me, nb);
    k_dist_repack<false><<<grid, 256, 0, st>>>(d, g1p, R, nb);
    return cudaGetLastError();
}

int dist_exchange(const DistHost& h, NcclComm comm, bool forward, cplx* S, cplx* R, int nb, cudaStream_t st,
                  const char** msg)
{
    const int me = h.me;
    const size_t nst_me = (size_t)h.nsw[me], nz_me = (size_t)h.nr3p[me];
    auto s_off = [&](int q) { return (size_t)h.i0r3p[q] * nb * nst_me; };       // my sticks, z range of rank q
    auto s_cnt = [&](int q) { return (size_t)h.nr3p[q] * nb * nst_me; };
    auto r_off = [&](int p) { return (size_t)nb * nz_me * h.prefix[p]; };       // rank p's sticks, my z range
    auto r_cnt = [&](int p) { return nz_me * nb * (size_t)h.nsw[p]; };
    // own block: device copy
    if (s_cnt(me)) {
        cudaError_t e = forward ? cudaMemcpyAsync(R + r_off(me), S + s_off(me), s_cnt(me) * 16, cudaMemcpyDeviceToDevice, st)
                                : cudaMemcpyAsync(S + s_off(me), R + r_off(me), s_cnt(me) * 16, cudaMemcpyDeviceToDevice, st);
        if (e != cudaSuccess) {
            *msg = cudaGetErrorString(e);
            return 1000 + (int)e;
        }
    }
    if (h.nproc == 1) return 0;
    const NcclApi& A = nccl_api();
    if (!A.ok || !comm) {
        *msg = A.ok ? "no NCCL communicator set (qeb_nccl_init / qeb_set_nccl)" : A.why;
        return 900;
    }
    int rc = A.GroupStart();
    for (int q = 0; q < h.nproc && rc == 0; ++q) {
        if (q == me) continue;
        if (forward) {
            if (s_cnt(q)) rc = A.Send(S + s_off(q), 2 * s_cnt(q), kNcclDouble, q, comm, st);
            if (rc == 0 && r_cnt(q)) rc = A.Recv(R + r_off(q), 2 * r_cnt(q), kNcclDouble, q, comm, st);
        } else {
            if (r_cnt(q)) rc = A.Send(R + r_off(q), 2 * r_cnt(q), kNcclDouble, q, comm, st);
            if (rc == 0 && s_cnt(q)) rc = A.Recv(S + s_off(q), 2 * s_cnt(q), kNcclDouble, q, comm, st);
        }
    }
    const int rc2 = A.GroupEnd();
    if (rc == 0) rc = rc2;
    if (rc != 0) {
        *msg = A.GetErrorString(rc);
        return 900 + rc;
    }
    return 0;
}

// after the puts of this exchange: flag word (dir, slot, me) of every peer = epoch
__global__ void k_p2p_signal(unsigned* const* __restrict__ peer_flags, int nproc, int me, int word0, unsigned epoch)
{
    const int q = threadIdx.x;
    if (q >= nproc || q == me) return;
    __threadfence_system();
    volatile unsigned* f = peer_flags[q] + word0 + me;
    *f = epoch;
    __threadfence_system();
}

// before the consumer kernels: wait until every peer has raised its word to >= epoch (their puts have landed).
// A peer that never arrives (crash, mismatched call sequence) traps after ~10 s instead of hanging the GPU.
__global__ void k_p2p_wait(const unsigned* __restrict__ flags, int nproc, int me, unsigned epoch)
{
    const int q = threadIdx.x;
    if (q >= nproc || q == me) return;
    const volatile unsigned* f = flags + q;
    const long long t0 = clock64();
    while ((int)(*f - epoch) < 0) {
        if (clock64() - t0 > 20000000000LL) __trap();
        __nanosleep(200);
    }
    __threadfence_system();
}

int dist_exchange_p2p(const DistHost& h, P2PState& p, bool forward, int slot, cplx* S, cplx* R, int nb, int nr3,
                      cudaStream_t st, const char** msg)
{
    const int me = h.me, np = h.nproc;
    const size_t nst_me = (size_t)h.nsw[me], nz_me = (size_t)h.nr3p[me];
    auto s_off = [&](int q) { return (size_t)h.i0r3p[q] * nb * nst_me; };
    auto s_cnt = [&](int q) { return (size_t)h.nr3p[q] * nb * nst_me; };
    auto r_off = [&](int q) { return (size_t)nb * nz_me * h.prefix[q]; };
    auto r_cnt = [&](int q) { return nz_me * nb * (size_t)h.nsw[q]; };
    cudaError_t e = cudaSuccess;
    if (s_cnt(me))
        e = forward ? cudaMemcpyAsync(R + r_off(me), S + s_off(me), s_cnt(me) * 16, cudaMemcpyDeviceToDevice, st)
                    : cudaMemcpyAsync(S + s_off(me), R + r_off(me), s_cnt(me) * 16, cudaMemcpyDeviceToDevice, st);
    for (int k = 1; k < np && e == cudaSuccess; ++k) {
        const int q = (me + k) % np;  // staggered: at step k every rank writes to a different peer
        if (forward) {
            if (!s_cnt(q)) continue;
            // peer q's R slice of this slot, block of rank me: [zl of q][box][my sticks]
            cplx* dst = p.peer_R[q] + (size_t)slot * ((size_t)h.nst_tot * h.nr3p[q] * p.batch) +
                        (size_t)nb * h.nr3p[q] * h.prefix[me];
            e = cudaMemcpyAsync(dst, S + s_off(q), s_cnt(q) * 16, cudaMemcpyDefault, st);
        } else {
            if (!r_cnt(q)) continue;
            // peer q's S slice of this slot, my z range of q's sticks: [z of me][box][q's sticks]
            cplx* dst = p.peer_S[q] + (size_t)slot * ((size_t)h.nsw[q] * nr3 * p.batch) +
                        (size_t)h.i0r3p[me] * nb * h.nsw[q];
            e = cudaMemcpyAsync(dst, R + r_off(q), r_cnt(q) * 16, cudaMemcpyDefault, st);
        }
    }
    if (e != cudaSuccess) {
        *msg = cudaGetErrorString(e);
        return 1000 + (int)e;
    }
    const int dir = forward ? 0 : 1;
    const unsigned epoch = ++p.epoch[dir][slot];
    const int word0 = (dir * p.nslots + slot) * np;
    const int thr = ((np + 31) / 32) * 32;
    k_p2p_signal<<<1, thr, 0, st>>>(p.peer_flags_dev, np, me, word0, epoch);
    k_p2p_wait<<<1, thr, 0, st>>>(p.peer_flags[me] + word0, np, me, epoch);
    e = cudaGetLastError();
    if (e != cudaSuccess) {
        *msg = cudaGetErrorString(e);
        return 1000 + (int)e;
    }
    return 0;
}

int dist_p2p_sync(const DistHost& h, P2PState& p, bool forward, int slot, cudaStream_t st, const char** msg)
{
    const int np = h.nproc, me = h.me;
    const int dir = forward ? 0 : 1;
    const unsigned epoch = ++p.epoch[dir][slot];
    const int word0 = (dir * p.nslots + slot) * np;
    const int thr = ((np + 31) / 32) * 32;
    k_p2p_signal<<<1, thr, 0, st>>>(p.peer_flags_dev, np, me, word0, epoch);
    k_p2p_wait<<<1, thr, 0, st>>>(p.peer_flags[me] + word0, np, me, epoch);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        *msg = cudaGetErrorString(e);
        return 1000 + (int)e;
    }
    return 0;
}

}  // namespace qeb
