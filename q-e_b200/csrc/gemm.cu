// Nonlocal projector contractions on the FP64 tensor pipe (DMMA, SASS
// DMMA.8x8x4 via mma.sync.m8n8k4.f64):
//   calbec      becp = vkb^H psi                  Modules/becmod.f90:431-486 (k), :274-334 (Gamma)
//   add_vuspsi  hpsi += vkb (blockdiag(D) becp)   PW/src/add_vuspsi.f90:125-171 (k), :68-122 (Gamma)
// Complex arithmetic is mapped onto the real MMA without de-interleaving:
//   calbec : the K index runs over the 2*npw real numbers of a column; the
//            imaginary part of the result uses the same A fragment and the B
//            fragment J*psi = (Im psi, -Re psi), formed in registers.
//   add_vu : computed transposed (MMA-M = bands, MMA-N = real rows of hpsi,
//            K = (projector, re/im)); the B fragment [vr -vi; vi vr] is formed
//            in registers from one 16-byte load of vkb(g,i).
// Both kernels: 256 threads, 8 warps, warp tile 32x(64 real columns), cp.async
// multi-stage pipeline, split-K with a fixed-order reduction (deterministic).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include "common.cuh"
#include "gemm.h"

namespace qeb {

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// sign flip on the integer pipe: an FP64 negation would be a DADD that competes with DMMA for the FP64 datapath
__device__ __forceinline__ double flip_sign(double x, unsigned mask)
{
    return __hiloint2double(__double2hiint(x) ^ (int)mask, __double2loint(x));
}

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc, bool valid)
{
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    const int sz = valid ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(d), "l"(gsrc), "r"(sz));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait()
{
    asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}

// ---------------------------------------------------------------------------
// calbec partial products
// ---------------------------------------------------------------------------
constexpr int CB_MT = 128;      // projectors per CTA
constexpr int CB_KC = 32;       // doubles of K per stage (16 complex G-vectors)
constexpr int CB_LDK = CB_KC + 4;
constexpr int CB_STAGES = 3;
template <bool CPLX>
struct CbCfg {
    static constexpr int NT = CPLX ? 64 : 128;  // bands per CTA
    static constexpr int NJ = CPLX ? 4 : 8;     // 8-band blocks per warp
    static constexpr size_t stage_doubles = (size_t)(CB_MT + NT) * CB_LDK;
    static constexpr size_t smem = stage_doubles * 8 * CB_STAGES;
};

template <bool CPLX>
__global__ void __launch_bounds__(256, 1)
k_calbec(const cplx* __restrict__ vkb, int npwx, const cplx* __restrict__ psi, int lda, int n, int nkb, int m,
         int gchunk, void* __restrict__ part)
{
    using C = CbCfg<CPLX>;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* sm = reinterpret_cast<double*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp & 3, wn = warp >> 2;
    const int r = lane >> 2, k = lane & 3;
    const int i0 = blockIdx.x * CB_MT, j0 = blockIdx.y * C::NT;
    const int g_beg = blockIdx.z * gchunk;
    const int g_end = min(n, g_beg + gchunk);
    const int nkt = (g_end - g_beg + 15) / 16;

    double acc[4][CPLX ? 8 : 8][2];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 8; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;

    auto load_stage = [&](int st, int kt) {
        double* As = sm + (size_t)st * C::stage_doubles;
        double* Bs = As + (size_t)CB_MT * CB_LDK;
        const int g0 = g_beg + kt * 16;
#pragma unroll
        for (int it = 0; it < CB_MT * 16 / 256; ++it) {
            const int c = tid + it * 256;
            const int row = c >> 4, cc = c & 15;
            const bool ok = (i0 + row < nkb) && (g0 + cc < g_end) && kt < nkt;
            const cplx* src = ok ? vkb + (size_t)(i0 + row) * npwx + g0 + cc : vkb;
            cp_async16(As + row * CB_LDK + 2 * cc, src, ok);
        }
#pragma unroll
        for (int it = 0; it < C::NT * 16 / 256; ++it) {
            const int c = tid + it * 256;
            const int row = c >> 4, cc = c & 15;
            const bool ok = (j0 + row < m) && (g0 + cc < g_end) && kt < nkt;
            const cplx* src = ok ? psi + (size_t)(j0 + row) * lda + g0 + cc : psi;
            cp_async16(Bs + row * CB_LDK + 2 * cc, src, ok);
        }
    };

#pragma unroll
    for (int s = 0; s < CB_STAGES - 1; ++s) {
        load_stage(s, s);
        cp_async_commit();
    }
    for (int kt = 0; kt < nkt; ++kt) {
        cp_async_wait<CB_STAGES - 2>();
        __syncthreads();
        load_stage((kt + CB_STAGES - 1) % CB_STAGES, kt + CB_STAGES - 1);
        cp_async_commit();
        const double* As = sm + (size_t)(kt % CB_STAGES) * C::stage_doubles;
        const double* Bs = As + (size_t)CB_MT * CB_LDK;
        const double* Aw = As + (size_t)(wm * 32 + r) * CB_LDK + k;
#pragma unroll
        for (int kk = 0; kk < CB_KC; kk += 4) {
            double a[4];
#pragma unroll
            for (int mi = 0; mi < 4; ++mi) a[mi] = Aw[(size_t)mi * 8 * CB_LDK + kk];
            if (CPLX) {
                // Re: B = psi_real[k]; Im: B = (J psi)[k] = (Im, -Re)[k&1], read from the neighbouring double
                const double* Bw = Bs + (size_t)(wn * 32 + r) * CB_LDK + (k & 2);
                const int sel = k & 1;
                const unsigned smask = sel ? 0x80000000u : 0u;
#pragma unroll
                for (int nj = 0; nj < 4; ++nj) {
                    const double bre = Bw[(size_t)nj * 8 * CB_LDK + kk + sel];
                    const double bim = flip_sign(Bw[(size_t)nj * 8 * CB_LDK + kk + (sel ^ 1)], smask);
#pragma unroll
                    for (int mi = 0; mi < 4; ++mi) {
                        dmma(acc[mi][nj][0], acc[mi][nj][1], a[mi], bre);
                        dmma(acc[mi][nj + 4][0], acc[mi][nj + 4][1], a[mi], bim);
                    }
                }
            } else {
                const double* Bw = Bs + (size_t)(wn * 64 + r) * CB_LDK + k;
#pragma unroll
                for (int nj = 0; nj < 8; ++nj) {
                    const double b = Bw[(size_t)nj * 8 * CB_LDK + kk];
#pragma unroll
                    for (int mi = 0; mi < 4; ++mi) dmma(acc[mi][nj][0], acc[mi][nj][1], a[mi], b);
                }
            }
        }
    }
    cp_async_wait<0>();
    // partials: part[s][j][i]
    const size_t sbase = (size_t)blockIdx.z * m;
#pragma unroll
    for (int mi = 0; mi < 4; ++mi) {
        const int i = i0 + wm * 32 + mi * 8 + r;
        if (i >= nkb) continue;
        if (CPLX) {
            cplx* P = reinterpret_cast<cplx*>(part);
#pragma unroll
            for (int nj = 0; nj < 4; ++nj)
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int j = j0 + wn * 32 + nj * 8 + 2 * k + e;
                    if (j < m) P[(sbase + j) * nkb + i] = mk(acc[mi][nj][e], acc[mi][nj + 4][e]);
                }
        } else {
            double* P = reinterpret_cast<double*>(part);
#pragma unroll
            for (int nj = 0; nj < 8; ++nj)
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int j = j0 + wn * 64 + nj * 8 + 2 * k + e;
                    if (j < m) P[(sbase + j) * nkb + i] = acc[mi][nj][e];
                }
        }
    }
}

// ---------------------------------------------------------------------------
// calbec, complex case, 3-multiplication form. With a = vkb(g,i), b = psi(g,j):
//   P1 = sum ar*br, P2 = sum ai*bi, P3 = sum (ar+ai)*(bi-br)
//   Re <a|b> = P1 + P2,   Im <a|b> = P3 + P1 - P2
// i.e. three real MMAs per complex multiply-add instead of four (25 % fewer DMMA). The MMA K index is the complex
// G-vector; each thread reads its fragment elements as one 16-byte complex and forms the sums in registers. The
// result differs from the 4-multiplication form only by rounding (normwise error ~ eps*sqrt(K)*|a||b|, far below
// the 1e-11 parity bound). CTA tile 128 projectors x 48 bands, warp tile 32 x 24, 8 warps (4 x 2).
// ---------------------------------------------------------------------------
constexpr int C3_MT = 128;
constexpr int C3_NT = 48;
constexpr int C3_KG = 16;        // complex G-vectors per stage
constexpr int C3_LD = 20;        // complex per smem row: 16 + 4 pad (row stride = 4 mod 8 sixteen-byte banks)
constexpr int C3_STAGES = 3;
constexpr size_t C3_STAGE_CPLX = (size_t)(C3_MT + C3_NT) * C3_LD;
constexpr size_t C3_SMEM = C3_STAGE_CPLX * 16 * C3_STAGES;

template <int MI>  // 8-row MMA blocks per warp: 4 -> 8 warps (4 x 2), 2 -> 16 warps (8 x 2)
__global__ void __launch_bounds__(1024 / MI, 1)
k_calbec3m(const cplx* __restrict__ vkb, int npwx, const cplx* __restrict__ psi, int lda, int n, int nkb, int m,
           int gchunk, cplx* __restrict__ part)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    cplx* sm = reinterpret_cast<cplx*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NTHR = 1024 / MI, WM = 16 / MI;  // warps along the projector dimension
    const int wm = warp % WM, wn = warp / WM;
    const int r = lane >> 2, k = lane & 3;
    const int i0 = blockIdx.x * C3_MT, j0 = blockIdx.y * C3_NT;
    const int g_beg = blockIdx.z * gchunk;
    const int g_end = min(n, g_beg + gchunk);
    const int nkt = (g_end - g_beg + C3_KG - 1) / C3_KG;
    // valid 8x8 blocks of this warp's tile (edge tiles of the projector / band dimensions)
    const int miv = max(0, min(MI, (nkb - (i0 + wm * 8 * MI) + 7) / 8));
    const int njv = max(0, min(3, (m - (j0 + wn * 24) + 7) / 8));
    const bool full = miv == MI && njv == 3;

    double p1[MI][3][2], p2[MI][3][2], p3[MI][3][2];
#pragma unroll
    for (int a = 0; a < MI; ++a)
#pragma unroll
        for (int b = 0; b < 3; ++b)
#pragma unroll
            for (int e = 0; e < 2; ++e) p1[a][b][e] = p2[a][b][e] = p3[a][b][e] = 0.0;

    auto load_stage = [&](int st, int kt) {
        cplx* As = sm + (size_t)st * C3_STAGE_CPLX;
        cplx* Bs = As + (size_t)C3_MT * C3_LD;
        const int g0 = g_beg + kt * C3_KG;
#pragma unroll
        for (int it = 0; it < C3_MT * C3_KG / NTHR; ++it) {
            const int c = tid + it * NTHR;
            const int row = c >> 4, cc = c & 15;
            const bool ok = (i0 + row < nkb) && (g0 + cc < g_end) && kt < nkt;
            const cplx* src = ok ? vkb + (size_t)(i0 + row) * npwx + g0 + cc : vkb;
            cp_async16(As + row * C3_LD + cc, src, ok);
        }
#pragma unroll
        for (int it = 0; it < (C3_NT * C3_KG + NTHR - 1) / NTHR; ++it) {
            const int c = tid + it * NTHR;
            if (c >= C3_NT * C3_KG) break;
            const int row = c >> 4, cc = c & 15;
            const bool ok = (j0 + row < m) && (g0 + cc < g_end) && kt < nkt;
            const cplx* src = ok ? psi + (size_t)(j0 + row) * lda + g0 + cc : psi;
            cp_async16(Bs + row * C3_LD + cc, src, ok);
        }
    };

#pragma unroll
    for (int s = 0; s < C3_STAGES - 1; ++s) {
        load_stage(s, s);
        cp_async_commit();
    }
    for (int kt = 0; kt < nkt; ++kt) {
        cp_async_wait<C3_STAGES - 2>();
        __syncthreads();
        load_stage((kt + C3_STAGES - 1) % C3_STAGES, kt + C3_STAGES - 1);
        cp_async_commit();
        const cplx* As = sm + (size_t)(kt % C3_STAGES) * C3_STAGE_CPLX;
        const cplx* Bs = As + (size_t)C3_MT * C3_LD;
        const cplx* Aw = As + (size_t)(wm * 8 * MI + r) * C3_LD + k;
        const cplx* Bw = Bs + (size_t)(wn * 24 + r) * C3_LD + k;
#pragma unroll
        for (int kk = 0; kk < C3_KG; kk += 4) {
            double ar[MI], ai[MI], as[MI];
#pragma unroll
            for (int mi = 0; mi < MI; ++mi) {
                const cplx a = Aw[(size_t)mi * 8 * C3_LD + kk];
                ar[mi] = a.x;
                ai[mi] = a.y;
                as[mi] = a.x + a.y;
            }
            // the sums are issued first and consumed last (P3), so the DMMAs of P1/P2 cover their latency
            double br[3], bi[3], bd[3];
#pragma unroll
            for (int nj = 0; nj < 3; ++nj) {
                const cplx b = Bw[(size_t)nj * 8 * C3_LD + kk];
                br[nj] = b.x;
                bi[nj] = b.y;
                bd[nj] = b.y - b.x;
            }
            if (full) {
#pragma unroll
                for (int nj = 0; nj < 3; ++nj)
#pragma unroll
                    for (int mi = 0; mi < MI; ++mi) {
                        dmma(p1[mi][nj][0], p1[mi][nj][1], ar[mi], br[nj]);
                        dmma(p2[mi][nj][0], p2[mi][nj][1], ai[mi], bi[nj]);
                    }
#pragma unroll
                for (int nj = 0; nj < 3; ++nj)
#pragma unroll
                    for (int mi = 0; mi < MI; ++mi) dmma(p3[mi][nj][0], p3[mi][nj][1], as[mi], bd[nj]);
            } else {  // edge tile: 8x8 blocks without a valid projector row or band column are skipped (warp-uniform)
#pragma unroll
                for (int nj = 0; nj < 3; ++nj)
#pragma unroll
                    for (int mi = 0; mi < MI; ++mi)
                        if (nj < njv && mi < miv) {
                            dmma(p1[mi][nj][0], p1[mi][nj][1], ar[mi], br[nj]);
                            dmma(p2[mi][nj][0], p2[mi][nj][1], ai[mi], bi[nj]);
                            dmma(p3[mi][nj][0], p3[mi][nj][1], as[mi], bd[nj]);
                        }
            }
        }
    }
    cp_async_wait<0>();
    // partials: part[s][j][i]
    const size_t sbase = (size_t)blockIdx.z * m;
#pragma unroll
    for (int mi = 0; mi < MI; ++mi) {
        const int i = i0 + wm * 8 * MI + mi * 8 + r;
        if (i >= nkb) continue;
#pragma unroll
        for (int nj = 0; nj < 3; ++nj)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int j = j0 + wn * 24 + nj * 8 + 2 * k + e;
                if (j < m)
                    part[(sbase + j) * nkb + i] =
                        mk(p1[mi][nj][e] + p2[mi][nj][e], p3[mi][nj][e] + (p1[mi][nj][e] - p2[mi][nj][e]));
            }
    }
}

// fixed-order reduction over the K splits (+ Gamma correction, becmod.f90:315-324)
template <bool CPLX>
__global__ void k_calbec_reduce(const void* __restrict__ part, int nsplit, int nkb, int m, void* __restrict__ becp,
                                const cplx* __restrict__ vkb, int npwx, const cplx* __restrict__ psi, int lda,
                                int gstart)
{
    const size_t tot = (size_t)nkb * m;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < tot; e += (size_t)gridDim.x * blockDim.x) {
        if (CPLX) {
            const cplx* P = reinterpret_cast<const cplx*>(part);
            double sx = 0.0, sy = 0.0;
            for (int s = 0; s < nsplit; ++s) {
                const cplx v = P[(size_t)s * tot + e];
                sx += v.x;
                sy += v.y;
            }
            reinterpret_cast<cplx*>(becp)[e] = mk(sx, sy);
        } else {
            const double* P = reinterpret_cast<const double*>(part);
            double sx = 0.0;
            for (int s = 0; s < nsplit; ++s) sx += P[(size_t)s * tot + e];
            sx *= 2.0;
            if (gstart == 2) {
                const int j = (int)(e / nkb), i = (int)(e - (size_t)j * nkb);
                sx -= vkb[(size_t)i * npwx].x * psi[(size_t)j * lda].x;
            }
            reinterpret_cast<double*>(becp)[e] = sx;
        }
    }
}

// ps(ofs+ih, j) = sum_jh deeq(ih,jh,na) becp(ofs+jh, j)   (add_vuspsi.f90:151-155, :92-97)
template <bool CPLX>
__global__ void k_apply_deeq(const void* __restrict__ becp, int nkb, int m, const int* __restrict__ row_atom,
                             const int* __restrict__ ofs, const int* __restrict__ nh, const double* __restrict__ deeq,
                             int nhm, void* __restrict__ ps, int ldps)
{
    const size_t tot = (size_t)nkb * m;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < tot; e += (size_t)gridDim.x * blockDim.x) {
        const int j = (int)(e / nkb), i = (int)(e - (size_t)j * nkb);
        const int na = row_atom[i];
        if (na < 0) {  // row not covered by any atom: contributes nothing
            if (CPLX)
                reinterpret_cast<cplx*>(ps)[(size_t)j * ldps + i] = mk(0.0, 0.0);
            else
                reinterpret_cast<double*>(ps)[(size_t)j * ldps + i] = 0.0;
            continue;
        }
        const int o = ofs[na], ih = i - o, nn = nh[na];
        const double* D = deeq + (size_t)na * nhm * nhm + ih;
        if (CPLX) {
            const cplx* B = reinterpret_cast<const cplx*>(becp) + (size_t)j * nkb + o;
            double sx = 0.0, sy = 0.0;
            for (int jh = 0; jh < nn; ++jh) {
                const double d = D[(size_t)jh * nhm];
                sx += d * B[jh].x;
                sy += d * B[jh].y;
            }
            reinterpret_cast<cplx*>(ps)[(size_t)j * ldps + i] = mk(sx, sy);
        } else {
            const double* B = reinterpret_cast<const double*>(becp) + (size_t)j * nkb + o;
            double sx = 0.0;
            for (int jh = 0; jh < nn; ++jh) sx += D[(size_t)jh * nhm] * B[jh];
            reinterpret_cast<double*>(ps)[(size_t)j * ldps + i] = sx;
        }
    }
}

// noncollinear: ps(ofs+ih, ipol, b) = sum_jh deeq_nc(ih,jh,na,1+2 ipol) becp(ofs+jh,1,b) + deeq_nc(ih,jh,na,2+2 ipol) becp(ofs+jh,2,b)
// (add_vuspsi.f90:192-223). becp / ps columns: 2b = up, 2b+1 = down. deeq_nc complex, Fortran (nhm,nhm,nat,4).
__global__ void k_apply_deeq_nc(const cplx* __restrict__ becp, int nkb, int m, const int* __restrict__ row_atom,
                                const int* __restrict__ ofs, const int* __restrict__ nh, const cplx* __restrict__ dnc,
                                int nhm, int nat, cplx* __restrict__ ps, int ldps)
{
    const size_t tot = (size_t)nkb * 2 * m;
    const size_t dspin = (size_t)nhm * nhm * nat;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < tot; e += (size_t)gridDim.x * blockDim.x) {
        const int col = (int)(e / nkb), i = (int)(e - (size_t)col * nkb);
        const int b = col >> 1, ipol = col & 1;
        const int na = row_atom[i];
        if (na < 0) {
            ps[(size_t)col * ldps + i] = mk(0.0, 0.0);
            continue;
        }
        const int o = ofs[na], ih = i - o, nn = nh[na];
        const cplx* D1 = dnc + (size_t)(2 * ipol) * dspin + (size_t)na * nhm * nhm + ih;
        const cplx* D2 = D1 + dspin;
        const cplx* B1 = becp + (size_t)(2 * b) * nkb + o;
        const cplx* B2 = B1 + nkb;
        double sx = 0.0, sy = 0.0;
        for (int jh = 0; jh < nn; ++jh) {
            const cplx d1 = D1[(size_t)jh * nhm], d2 = D2[(size_t)jh * nhm];
            const cplx b1 = B1[jh], b2 = B2[jh];
            sx += d1.x * b1.x - d1.y * b1.y + d2.x * b2.x - d2.y * b2.y;
            sy += d1.x * b1.y + d1.y * b1.x + d2.x * b2.y + d2.y * b2.x;
        }
        ps[(size_t)col * ldps + i] = mk(sx, sy);
    }
}

// ---------------------------------------------------------------------------
// add_vuspsi: hpsi(g,j) += sum_i vkb(g,i) ps(i,j)
// ---------------------------------------------------------------------------
constexpr int AV_BT = 64;    // bands per CTA
constexpr int AV_GT = 128;   // G-vectors per CTA
constexpr int AV_LDV = AV_GT + 2;
constexpr int AV_KI = 16;    // projectors per stage
constexpr int AV_STAGES = 4;
template <bool CPLX>
struct AvCfg {
    static constexpr int KD = CPLX ? 32 : 16;  // doubles of ps per band per stage
    static constexpr int LDK = KD + 4;
    static constexpr size_t ps_doubles = (size_t)AV_BT * LDK;
    static constexpr size_t stage_doubles = ps_doubles + (size_t)AV_KI * AV_LDV * 2;
    static constexpr size_t smem = stage_doubles * 8 * AV_STAGES;
};

template <bool CPLX>
__global__ void __launch_bounds__(256, 1)
k_addvus(const cplx* __restrict__ vkb, int npwx, const void* __restrict__ ps_, int ldps, int n, int nkb, int m,
         cplx* __restrict__ hpsi, int lda)
{
    using C = AvCfg<CPLX>;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* sm = reinterpret_cast<double*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp & 1, wn = warp >> 1;
    const int r = lane >> 2, k = lane & 3;
    const int j0 = blockIdx.x * AV_BT, g0 = blockIdx.y * AV_GT;
    const int nkt = (nkb + AV_KI - 1) / AV_KI;
    const double* psd = reinterpret_cast<const double*>(ps_);
    const int ldps_d = CPLX ? 2 * ldps : ldps;  // doubles per band column

    double acc[4][8][2];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 8; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;

    auto load_stage = [&](int st, int kt) {
        double* Ps = sm + (size_t)st * C::stage_doubles;
        cplx* Vs = reinterpret_cast<cplx*>(Ps + C::ps_doubles);
        const int ib = kt * AV_KI;
        // ps tile: AV_BT bands x KD doubles, 16-byte chunks
        constexpr int CH = C::KD / 2;  // chunks per band
#pragma unroll
        for (int it = 0; it < (AV_BT * CH + 255) / 256; ++it) {
            const int c = tid + it * 256;
            if (c < AV_BT * CH) {
                const int row = c / CH, cc = c % CH;
                const int d0 = (CPLX ? 2 * ib : ib) + 2 * cc;  // first double index within the band column
                const bool ok = (j0 + row < m) && (d0 < (CPLX ? 2 * nkb : ldps)) && kt < nkt;
                const double* src = ok ? psd + (size_t)(j0 + row) * ldps_d + d0 : psd;
                cp_async16(Ps + row * C::LDK + 2 * cc, src, ok);
            }
        }
        // vkb tile: AV_KI projectors x AV_GT G-vectors
#pragma unroll
        for (int it = 0; it < AV_KI * AV_GT / 256; ++it) {
            const int c = tid + it * 256;
            const int row = c / AV_GT, cc = c % AV_GT;
            const bool ok = (ib + row < nkb) && (g0 + cc < n) && kt < nkt;
            const cplx* src = ok ? vkb + (size_t)(ib + row) * npwx + g0 + cc : vkb;
            cp_async16(Vs + row * AV_LDV + cc, src, ok);
        }
    };

#pragma unroll
    for (int s = 0; s < AV_STAGES - 1; ++s) {
        load_stage(s, s);
        cp_async_commit();
    }
    for (int kt = 0; kt < nkt; ++kt) {
        cp_async_wait<AV_STAGES - 2>();
        __syncthreads();
        load_stage((kt + AV_STAGES - 1) % AV_STAGES, kt + AV_STAGES - 1);
        cp_async_commit();
        const double* Ps = sm + (size_t)(kt % AV_STAGES) * C::stage_doubles;
        const cplx* Vs = reinterpret_cast<const cplx*>(Ps + C::ps_doubles);
        const double* Aw = Ps + (size_t)(wm * 32 + r) * C::LDK + k;
#pragma unroll
        for (int kk = 0; kk < C::KD; kk += 4) {
            double a[4];
#pragma unroll
            for (int mi = 0; mi < 4; ++mi) a[mi] = Aw[(size_t)mi * 8 * C::LDK + kk];
            if (CPLX) {
                // K index (i, c'): i = kk/2 + (k>>1), c' = k&1 ; N index (g, c): g = .. + (r>>1), c = r&1
                // [vr -vi; vi vr]: element (c = r&1, c' = k&1) is the double at offset c^c', negated for (c=0,c'=1)
                const double* Vw = reinterpret_cast<const double*>(Vs + (size_t)(kk / 2 + (k >> 1)) * AV_LDV + wn * 32 + (r >> 1)) + ((r ^ k) & 1);
                const unsigned smask = (!(r & 1) && (k & 1)) ? 0x80000000u : 0u;
#pragma unroll
                for (int nj = 0; nj < 8; ++nj) {
                    const double b = flip_sign(Vw[nj * 8], smask);
#pragma unroll
                    for (int mi = 0; mi < 4; ++mi) dmma(acc[mi][nj][0], acc[mi][nj][1], a[mi], b);
                }
            } else {
                // K index i = kk + k ; N index: real row 2*g + c = 2*(wn*32 + nj*4) + r
                const double* Vw = reinterpret_cast<const double*>(Vs + (size_t)(kk + k) * AV_LDV + wn * 32) + r;
#pragma unroll
                for (int nj = 0; nj < 8; ++nj) {
                    const double b = Vw[nj * 8];
#pragma unroll
                    for (int mi = 0; mi < 4; ++mi) dmma(acc[mi][nj][0], acc[mi][nj][1], a[mi], b);
                }
            }
        }
    }
    cp_async_wait<0>();
#pragma unroll
    for (int mi = 0; mi < 4; ++mi) {
        const int j = j0 + wm * 32 + mi * 8 + r;
        if (j >= m) continue;
#pragma unroll
        for (int nj = 0; nj < 8; ++nj) {
            const int g = g0 + wn * 32 + nj * 4 + k;
            if (g < n) {
                cplx* h = hpsi + (size_t)j * lda + g;
                const cplx o = *h;
                *h = mk(o.x + acc[mi][nj][0], o.y + acc[mi][nj][1]);
            }
        }
    }
}

// ---------------------------------------------------------------------------
// add_vuspsi, complex case, 3-multiplication form. With p = ps(i,j), v = vkb(g,i):
//   P1 = sum pr*vr, P2 = sum pi*vi, P3 = sum (pr+pi)*(vr+vi)
//   Re (v p) = P1 - P2,   Im (v p) = P3 - P1 - P2
// MMA-M = bands, MMA-N = G-vectors, MMA-K = projectors (all complex indices; fragments are 16-byte complex loads).
// CTA tile 64 bands x 96 G-vectors, warp tile 32 x 24, 8 warps (2 x 4).
// ---------------------------------------------------------------------------
constexpr int A3_BT = 64;
constexpr int A3_GT = 96;
constexpr int A3_KI = 16;        // projectors per stage
constexpr int A3_LDP = 20;       // complex per ps row (16 + 4 pad)
constexpr int A3_LDV = 98;       // complex per vkb row (96 + 2 pad: row stride = 2 mod 8 sixteen-byte banks)
constexpr int A3_STAGES = 4;
constexpr size_t A3_STAGE_CPLX = (size_t)A3_BT * A3_LDP + (size_t)A3_KI * A3_LDV;
constexpr size_t A3_SMEM = A3_STAGE_CPLX * 16 * A3_STAGES;

template <int MI>  // 4 -> 8 warps (2 x 4), 2 -> 16 warps (4 x 4)
__global__ void __launch_bounds__(1024 / MI, 1)
k_addvus3m(const cplx* __restrict__ vkb, int npwx, const cplx* __restrict__ ps, int ldps, int n, int nkb, int m,
           cplx* __restrict__ hpsi, int lda)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    cplx* sm = reinterpret_cast<cplx*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NTHR = 1024 / MI, WM = 8 / MI;  // warps along the band dimension
    // consecutive warps (= the four SM sub-partitions) share one band slice: when the last band tile is only partly
    // populated, the warps that still have bands are spread over all four tensor units and the others skip their MMAs
    const int wm = warp / 4, wn = warp % 4;
    const int r = lane >> 2, k = lane & 3;
    const int j0 = blockIdx.x * A3_BT, g0 = blockIdx.y * A3_GT;
    const bool has_bands = (j0 + wm * 8 * MI) < m;
    const int nkt = (nkb + A3_KI - 1) / A3_KI;

    double p1[MI][3][2], p2[MI][3][2], p3[MI][3][2];
#pragma unroll
    for (int a = 0; a < MI; ++a)
#pragma unroll
        for (int b = 0; b < 3; ++b)
#pragma unroll
            for (int e = 0; e < 2; ++e) p1[a][b][e] = p2[a][b][e] = p3[a][b][e] = 0.0;

    auto load_stage = [&](int st, int kt) {
        cplx* Ps = sm + (size_t)st * A3_STAGE_CPLX;
        cplx* Vs = Ps + (size_t)A3_BT * A3_LDP;
        const int ib = kt * A3_KI;
        // ps tile: 64 bands x 16 projectors
#pragma unroll
        for (int it = 0; it < A3_BT * A3_KI / NTHR; ++it) {
            const int c = tid + it * NTHR;
            const int row = c >> 4, cc = c & 15;
            const bool ok = (j0 + row < m) && (ib + cc < nkb) && kt < nkt;
            const cplx* src = ok ? ps + (size_t)(j0 + row) * ldps + ib + cc : ps;
            cp_async16(Ps + row * A3_LDP + cc, src, ok);
        }
        // vkb tile: 16 projectors x 96 G-vectors
#pragma unroll
        for (int it = 0; it < A3_KI * A3_GT / NTHR; ++it) {
            const int c = tid + it * NTHR;
            const int row = c / A3_GT, cc = c - row * A3_GT;
            const bool ok = (ib + row < nkb) && (g0 + cc < n) && kt < nkt;
            const cplx* src = ok ? vkb + (size_t)(ib + row) * npwx + g0 + cc : vkb;
            cp_async16(Vs + row * A3_LDV + cc, src, ok);
        }
    };

#pragma unroll
    for (int s = 0; s < A3_STAGES - 1; ++s) {
        load_stage(s, s);
        cp_async_commit();
    }
    for (int kt = 0; kt < nkt; ++kt) {
        cp_async_wait<A3_STAGES - 2>();
        __syncthreads();
        load_stage((kt + A3_STAGES - 1) % A3_STAGES, kt + A3_STAGES - 1);
        cp_async_commit();
        const cplx* Ps = sm + (size_t)(kt % A3_STAGES) * A3_STAGE_CPLX;
        const cplx* Vs = Ps + (size_t)A3_BT * A3_LDP;
        const cplx* Aw = Ps + (size_t)(wm * 8 * MI + r) * A3_LDP + k;   // A[band][i]
        const cplx* Bw = Vs + (size_t)k * A3_LDV + wn * 24 + r;     // B[i][g]
        if (has_bands)
#pragma unroll
        for (int kk = 0; kk < A3_KI; kk += 4) {
            double ar[MI], ai[MI], as[MI];
#pragma unroll
            for (int mi = 0; mi < MI; ++mi) {
                const cplx a = Aw[(size_t)mi * 8 * A3_LDP + kk];
                ar[mi] = a.x;
                ai[mi] = a.y;
                as[mi] = a.x + a.y;
            }
            double br[3], bi[3], bs[3];
#pragma unroll
            for (int nj = 0; nj < 3; ++nj) {
                const cplx b = Bw[(size_t)kk * A3_LDV + nj * 8];
                br[nj] = b.x;
                bi[nj] = b.y;
                bs[nj] = b.x + b.y;
            }
#pragma unroll
            for (int nj = 0; nj < 3; ++nj)
#pragma unroll
                for (int mi = 0; mi < MI; ++mi) {
                    dmma(p1[mi][nj][0], p1[mi][nj][1], ar[mi], br[nj]);
                    dmma(p2[mi][nj][0], p2[mi][nj][1], ai[mi], bi[nj]);
                }
#pragma unroll
            for (int nj = 0; nj < 3; ++nj)
#pragma unroll
                for (int mi = 0; mi < MI; ++mi) dmma(p3[mi][nj][0], p3[mi][nj][1], as[mi], bs[nj]);
        }
    }
    cp_async_wait<0>();
#pragma unroll
    for (int mi = 0; mi < MI; ++mi) {
        const int j = j0 + wm * 8 * MI + mi * 8 + r;
        if (j >= m) continue;
#pragma unroll
        for (int nj = 0; nj < 3; ++nj)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int g = g0 + wn * 24 + nj * 8 + 2 * k + e;
                if (g < n) {
                    cplx* h = hpsi + (size_t)j * lda + g;
                    const cplx o = *h;
                    const double re = p1[mi][nj][e] - p2[mi][nj][e];
                    const double im = p3[mi][nj][e] - (p1[mi][nj][e] + p2[mi][nj][e]);
                    *h = mk(o.x + re, o.y + im);
                }
            }
    }
}

// Gamma: Im hpsi(G=0) = 0 (h_psi.f90:294-295)
__global__ void k_zero_im_g0(cplx* hpsi, int lda, int m)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < m) hpsi[(size_t)j * lda].y = 0.0;
}

// ---------------------------------------------------------------------------
// host launchers
// ---------------------------------------------------------------------------
static int g_smem_set = 0;
static int g_3m_warps = 8;  // measured on B200: 8 warps (warp tile 32x24) 63.3 ms, 16 warps (16x24) 65.2 ms per AUSURF112 call
static bool g_use_3m = true;  // complex projector GEMMs in the 3-multiplication form (QEB_GEMM_4M=1: the 4-mult kernels)
static cudaError_t set_smem_attrs()
{
    if (g_smem_set) return cudaSuccess;
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(k_calbec<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CbCfg<true>::smem))) return e;
    if ((e = cudaFuncSetAttribute(k_calbec<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CbCfg<false>::smem))) return e;
    if ((e = cudaFuncSetAttribute(k_addvus<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)AvCfg<true>::smem))) return e;
    if ((e = cudaFuncSetAttribute(k_addvus<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)AvCfg<false>::smem))) return e;
    if ((e = cudaFuncSetAttribute(k_calbec3m<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C3_SMEM))) return e;
    if ((e = cudaFuncSetAttribute(k_addvus3m<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)A3_SMEM))) return e;
    if ((e = cudaFuncSetAttribute(k_calbec3m<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C3_SMEM))) return e;
    if ((e = cudaFuncSetAttribute(k_addvus3m<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)A3_SMEM))) return e;
    if (const char* v = getenv("QEB_GEMM_WARPS")) g_3m_warps = atoi(v) == 16 ? 16 : 8;
    if (const char* v = getenv("QEB_GEMM_4M")) g_use_3m = atoi(v) == 0;
    g_smem_set = 1;
    return cudaSuccess;
}

void calbec_plan(int n, int nkb, int m, bool cplx_mode, int num_sms, int* nsplit, int* gchunk)
{
    const int NT = cplx_mode ? (g_use_3m ? C3_NT : 64) : 128;
    const long tiles = (long)((nkb + CB_MT - 1) / CB_MT) * ((m + NT - 1) / NT);
    const int max_split = 64;
    // each split handles >= 256 G-vectors (16 k-tiles); pick the split count
    // that best fills whole waves of num_sms CTAs
    int smax = n / 256;
    if (smax < 1) smax = 1;
    if (smax > max_split) smax = max_split;
    int best = 1;
    double beste = -1.0;
    for (int s = 1; s <= smax; ++s) {
        const long ctas = tiles * s;
        const long waves = (ctas + num_sms - 1) / num_sms;
        double eff = (double)ctas / (double)(waves * num_sms);
        // mild preference for fewer splits (less reduction traffic)
        eff -= 0.002 * s;
        if (eff > beste) {
            beste = eff;
            best = s;
        }
    }
    int chunk = (n + best - 1) / best;
    chunk = (chunk + 15) / 16 * 16;
    if (chunk < 16) chunk = 16;
    best = (n + chunk - 1) / chunk;
    if (best < 1) best = 1;
    *nsplit = best;
    *gchunk = chunk;
}

size_t calbec_workspace_bytes(int nkb, int m, int nsplit, bool cplx_mode)
{
    return (size_t)nsplit * nkb * m * (cplx_mode ? 16 : 8);
}

cudaError_t launch_calbec(const cplx* vkb, int npwx, const cplx* psi, int lda, int n, int nkb, int m, bool cplx_mode,
                          int gstart, int nsplit, int gchunk, void* part, void* becp, cudaStream_t st)
{
    cudaError_t e = set_smem_attrs();
    if (e) return e;
    if (nkb <= 0 || m <= 0) return cudaSuccess;
    const int NT = cplx_mode ? (g_use_3m ? C3_NT : 64) : 128;
    dim3 grid((nkb + CB_MT - 1) / CB_MT, (m + NT - 1) / NT, nsplit);
    if (cplx_mode && g_use_3m) {
        if (g_3m_warps == 16)
            k_calbec3m<2><<<grid, 512, C3_SMEM, st>>>(vkb, npwx, psi, lda, n, nkb, m, gchunk, (cplx*)part);
        else
            k_calbec3m<4><<<grid, 256, C3_SMEM, st>>>(vkb, npwx, psi, lda, n, nkb, m, gchunk, (cplx*)part);
    } else if (cplx_mode)
        k_calbec<true><<<grid, 256, CbCfg<true>::smem, st>>>(vkb, npwx, psi, lda, n, nkb, m, gchunk, part);
    else
        k_calbec<false><<<grid, 256, CbCfg<false>::smem, st>>>(vkb, npwx, psi, lda, n, nkb, m, gchunk, part);
    const size_t tot = (size_t)nkb * m;
    int blocks = (int)((tot + 255) / 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (cplx_mode)
        k_calbec_reduce<true><<<blocks, 256, 0, st>>>(part, nsplit, nkb, m, becp, vkb, npwx, psi, lda, gstart);
    else
        k_calbec_reduce<false><<<blocks, 256, 0, st>>>(part, nsplit, nkb, m, becp, vkb, npwx, psi, lda, gstart);
    return cudaGetLastError();
}

cudaError_t launch_apply_deeq(const void* becp, int nkb, int m, bool cplx_mode, const int* row_atom, const int* ofs,
                              const int* nh, const double* deeq, int nhm, void* ps, int ldps, cudaStream_t st)
{
    if (nkb <= 0 || m <= 0) return cudaSuccess;
    const size_t tot = (size_t)nkb * m;
    int blocks = (int)((tot + 255) / 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (cplx_mode)
        k_apply_deeq<true><<<blocks, 256, 0, st>>>(becp, nkb, m, row_atom, ofs, nh, deeq, nhm, ps, ldps);
    else
        k_apply_deeq<false><<<blocks, 256, 0, st>>>(becp, nkb, m, row_atom, ofs, nh, deeq, nhm, ps, ldps);
    return cudaGetLastError();
}

cudaError_t launch_apply_deeq_nc(const void* becp, int nkb, int m, const int* row_atom, const int* ofs, const int* nh,
                                 const void* deeq_nc, int nhm, int nat, void* ps, int ldps, cudaStream_t st)
{
    const size_t tot = (size_t)nkb * 2 * m;
    if (tot == 0) return cudaSuccess;
    int blocks = (int)((tot + 255) / 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    k_apply_deeq_nc<<<blocks, 256, 0, st>>>((const cplx*)becp, nkb, m, row_atom, ofs, nh, (const cplx*)deeq_nc, nhm, nat,
                                            (cplx*)ps, ldps);
    return cudaGetLastError();
}

cudaError_t launch_addvus(const cplx* vkb, int npwx, const void* ps, int ldps, int n, int nkb, int m, bool cplx_mode,
                          cplx* hpsi, int lda, cudaStream_t st)
{
    cudaError_t e = set_smem_attrs();
    if (e) return e;
    if (nkb <= 0 || m <= 0 || n <= 0) return cudaSuccess;
    dim3 grid((m + AV_BT - 1) / AV_BT, (n + AV_GT - 1) / AV_GT);
    if (cplx_mode && g_use_3m) {
        dim3 grid3((m + A3_BT - 1) / A3_BT, (n + A3_GT - 1) / A3_GT);
        if (g_3m_warps == 16)
            k_addvus3m<2><<<grid3, 512, A3_SMEM, st>>>(vkb, npwx, (const cplx*)ps, ldps, n, nkb, m, hpsi, lda);
        else
            k_addvus3m<4><<<grid3, 256, A3_SMEM, st>>>(vkb, npwx, (const cplx*)ps, ldps, n, nkb, m, hpsi, lda);
    } else if (cplx_mode)
        k_addvus<true><<<grid, 256, AvCfg<true>::smem, st>>>(vkb, npwx, ps, ldps, n, nkb, m, hpsi, lda);
    else
        k_addvus<false><<<grid, 256, AvCfg<false>::smem, st>>>(vkb, npwx, ps, ldps, n, nkb, m, hpsi, lda);
    return cudaGetLastError();
}

cudaError_t launch_zero_im_g0(cplx* hpsi, int lda, int m, cudaStream_t st)
{
    if (m <= 0) return cudaSuccess;
    k_zero_im_g0<<<(m + 127) / 128, 128, 0, st>>>(hpsi, lda, m);
    return cudaGetLastError();
}

}  // namespace qeb
