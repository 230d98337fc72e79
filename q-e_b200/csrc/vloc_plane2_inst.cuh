// plane2 kernel (vloc_plane2.cuh): kernel template + launcher, instantiated once per (x plan, y plan, threads) in
// vloc_plane2_<n1>_<n2>.cu so that the instantiations compile in parallel.
//
// The kernel is persistent: CTA c owns planes q in [Q c/G, Q (c+1)/G) of the flat list q = box*nr3 + z (Q = nr3*nboxes),
// so any batch size fills the SMs evenly. The stick row of plane q+1 is fetched with ONE bulk asynchronous copy
// (cp.async.bulk.shared::cluster.global + mbarrier complete_tx, SASS UBLKCP) issued by one thread right after the first y
// pass has consumed the current row, and the result row leaves with one cp.async.bulk.global.shared::cta; no thread
// spends issue slots on the stick traffic.
#pragma once
#include <cuda_runtime.h>
#include <stdlib.h>
#include "vloc.h"
#include "vloc_plane2.cuh"

namespace qeb {

QEB_D unsigned p2_saddr(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
QEB_D void p2_mbar_init(void* bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(p2_saddr(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
}
QEB_D void p2_bulk_load(void* smem_dst, const void* gsrc, unsigned bytes, void* bar)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(p2_saddr(bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(
                     p2_saddr(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(p2_saddr(bar))
                 : "memory");
}
QEB_D void p2_mbar_wait(void* bar, unsigned parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "P2_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra P2_DONE;\n"
        "bra P2_WAIT;\n"
        "P2_DONE:\n"
        "}\n" ::"r"(p2_saddr(bar)),
        "r"(parity)
        : "memory");
}
QEB_D void p2_bulk_store(void* gdst, const void* smem_src, unsigned bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n" ::"l"(gdst), "r"(p2_saddr(smem_src)), "r"(bytes)
                 : "memory");
    asm volatile("cp.async.bulk.commit_group;\n" ::: "memory");
}
QEB_D void p2_bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory"); }
QEB_D void p2_bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;\n" ::: "memory"); }
QEB_D void p2_fence_async() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }

struct P2Ctx {  // what the unrolled phase list needs besides the phase arguments
    const cplx* next_row;  // stick row of the next plane (nullptr: none)
    cplx* stin_w;
    void* bar;
    unsigned row_bytes;
};

template <class PLX, class PLY, int PH>
struct P2Run {
    static __device__ __forceinline__ void go(int tid, int nthr, int nxw, const P2Roles& ro, cplx* P, const cplx* stin,
                                              cplx* stout, const Plane2Sm& sm, const double* vz, const P2Ctx& cx)
    {
        constexpr int NPH = Plane2<PLX, PLY>::NPH;
        plane2_phase_c<PLX, PLY, PH>(tid, nthr, nxw, ro, P, stin, stout, sm, vz);
        if constexpr (PH + 1 < NPH) {
            // the last phase rewrites stout: the bulk store of the previous plane must have read it
            if (PH + 2 == NPH && tid == 0) p2_bulk_wait_read();
            __syncthreads();
            // stin is consumed by phase 0: fetch the next plane's row behind the remaining phases
            if (PH == 0 && tid == 0 && cx.next_row) p2_bulk_load(cx.stin_w, cx.next_row, cx.row_bytes, cx.bar);
            P2Run<PLX, PLY, PH + 1>::go(tid, nthr, nxw, ro, P, stin, stout, sm, vz, cx);
        }
    }
};

template <class PLX, class PLY, int NT>
__global__ void __launch_bounds__(NT, 1) k_vloc_plane2(const __grid_constant__ VlocTables t, int nboxes)
{
    using G = Plane2<PLX, PLY>;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    cplx* P = reinterpret_cast<cplx*>(smem_raw);
    const int nst = t.nst, nxw = t.nxw;
    const int nstp = plane2_row_elems(nst);
    cplx* stin = P + (size_t)G::NY * G::LDP;
    cplx* stout = stin + nstp;
    cplx* twx = stout + nstp;
    cplx* twy = twx + G::NX;
    int* ymapT = reinterpret_cast<int*>(twy + G::NY);
    int* xactS = ymapT + (size_t)G::NY * nxw;
    int* xmaskS = xactS + nxw;
    unsigned long long* bar =
        reinterpret_cast<unsigned long long*>((reinterpret_cast<uintptr_t>(xmaskS + G::XMASK) + 15) & ~(uintptr_t)15);
    const int tid = threadIdx.x, nthr = blockDim.x;
    const long long Q = (long long)t.nr3 * nboxes;
    const int q0 = (int)(Q * blockIdx.x / gridDim.x), q1 = (int)(Q * (blockIdx.x + 1) / gridDim.x);
    if (q0 >= q1) return;
    int z = q0 % t.nr3, box = q0 / t.nr3;
    const unsigned row_bytes = (unsigned)nst * 16u;
    if (tid == 0) p2_mbar_init(bar, 1);
    __syncthreads();
    if (tid == 0) p2_bulk_load(stin, t.g1 + (size_t)box * t.g1_box + (size_t)z * t.g1_zs, row_bytes, bar);
    for (int i = tid; i < G::NX; i += nthr) twx[i] = t.px.tw[i];
    for (int i = tid; i < G::NY; i += nthr) twy[i] = t.py.tw[i];
    for (int i = tid; i < G::NY * nxw; i += nthr) ymapT[i] = plane2_ymap(t.pl_ymapT[i], nst);
    if (tid == 0) stin[nst] = mk(0.0, 0.0);  // the zero every stick-less column position reads (bulk loads stop before it)
    for (int i = tid; i < nxw; i += nthr) xactS[i] = t.pl_xact[i];
    if (tid < G::XMASK) {
        constexpr int R = PLX::r(0), A = G::XMASK;
        int m = 0;
#pragma unroll
        for (int b = 0; b < R; ++b)
            if (t.xmap[tid + A * b] >= 0) m |= 1 << b;
        xmaskS[tid] = m;
    }
    const Plane2Sm sm{twx, twy, ymapT, xactS, xmaskS};
    const P2Roles ro = plane2_roles(tid, nthr, nxw, G::NGX);
    unsigned parity = 0;
    __syncthreads();
#pragma unroll 1
    for (int q = q0; q < q1; ++q) {
        int zn = z + 1, bn = box;
        if (zn == t.nr3) {
            zn = 0;
            ++bn;
        }
        P2Ctx cx;
        cx.next_row = q + 1 < q1 ? t.g1 + (size_t)bn * t.g1_box + (size_t)zn * t.g1_zs : nullptr;
        cx.stin_w = stin;
        cx.bar = bar;
        cx.row_bytes = row_bytes;
        p2_mbar_wait(bar, parity);
        parity ^= 1u;
        P2Run<PLX, PLY, 0>::go(tid, nthr, nxw, ro, P, stin, stout, sm, t.pl_v + (size_t)z * (G::NX * G::NY), cx);
        p2_fence_async();  // the generic-proxy writes of stout become visible to the bulk store
        __syncthreads();
        if (tid == 0) p2_bulk_store(t.g1 + (size_t)box * t.g1_box + (size_t)z * t.g1_zs, stout, row_bytes);
        z = zn;
        box = bn;
    }
    if (tid == 0) p2_bulk_wait_all();
}

template <class PLX, class PLY, int NT>
struct Plane2Launch {
    static cudaError_t attrs(int s);
    // false: this instantiation does not match the tables' plans, or the plane does not fit
    static bool run(const VlocTables& t, int nboxes, int num_sms, size_t max_smem, cudaStream_t st);
};

#ifdef QEB_PLANE2_INSTANTIATE
template <class PLX, class PLY, int NT>
cudaError_t Plane2Launch<PLX, PLY, NT>::attrs(int s)
{
    return cudaFuncSetAttribute(k_vloc_plane2<PLX, PLY, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, s);
}
template <class PL>
static int p2_max_nj()
{
    int m = 1;
    for (int p = 0; p < PL::NP; ++p) m = (PL::N / PL::r(p)) > m ? (PL::N / PL::r(p)) : m;
    return m;
}
template <class PLX, class PLY, int NT>
bool Plane2Launch<PLX, PLY, NT>::run(const VlocTables& t, int nboxes, int num_sms, size_t max_smem, cudaStream_t st)
{
    using G = Plane2<PLX, PLY>;
    if (!PLX::matches(t.px) || !PLY::matches(t.py) || t.nst < 1) return false;
    const size_t smem = plane2_smem_bytes<PLX, PLY>(t.nst, t.nxw);
    if (smem > max_smem) return false;
    // warps: one per task of the largest phase, at most NT/32
    int tasks = G::NGX * p2_max_nj<PLX>();
    const int ty = ((t.nxw + 31) / 32) * p2_max_nj<PLY>();
    if (ty > tasks) tasks = ty;
    int nw = tasks < NT / 32 ? tasks : NT / 32;
    if (const char* e = getenv("QEB_NWP")) nw = atoi(e);
    if (nw < 1) nw = 1;
    if (nw * 32 > NT) nw = NT / 32;
    int per_sm = 1;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_vloc_plane2<PLX, PLY, NT>, nw * 32, smem) != cudaSuccess ||
        per_sm < 1)
        per_sm = 1;
    const long long Q = (long long)t.nr3 * nboxes;
    long long grid = (long long)num_sms * per_sm;
    if (grid > Q) grid = Q;
    k_vloc_plane2<PLX, PLY, NT><<<(unsigned)grid, nw * 32, smem, st>>>(t, nboxes);
    return true;
}
#endif

}  // namespace qeb
