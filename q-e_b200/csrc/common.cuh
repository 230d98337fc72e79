// Shared definitions for the qeb200 kernels (sm_100a) and for the host-side
// emulation harness used by the CPU tests (tests/emul): every kernel body is a
// sequence of barrier-separated phases written as host+device functions of
// (tid, nthreads), so the index/twiddle logic can be executed serially on a
// CPU with plain g++ while the product path only ever runs the CUDA kernels.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define QEB_HD __host__ __device__ __forceinline__
#define QEB_D __device__ __forceinline__
#else
#define QEB_HD inline
#define QEB_D inline
#endif

namespace qeb {

struct alignas(16) cplx {
    double x, y;
};

QEB_HD cplx mk(double x, double y) { cplx r; r.x = x; r.y = y; return r; }
QEB_HD cplx operator+(cplx a, cplx b) { return mk(a.x + b.x, a.y + b.y); }
QEB_HD cplx operator-(cplx a, cplx b) { return mk(a.x - b.x, a.y - b.y); }
QEB_HD cplx cmul(cplx a, cplx b) { return mk(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
QEB_HD cplx scal(cplx a, double s) { return mk(a.x * s, a.y * s); }
// multiply by +i (s=+1) or -i (s=-1)
QEB_HD cplx muli(cplx a, double s) { return mk(-s * a.y, s * a.x); }

// Division by a runtime constant d (1 <= d < 2^31) of x < 2^31 with a
// multiply-high: q = (umulhi(x, mul) + x) >> sh  (round-up method).
struct FastDiv {
    uint32_t d, mul, sh;
};
inline FastDiv make_fastdiv(uint32_t d)
{
    FastDiv f;
    f.d = d;
    uint32_t sh = 0;
    while ((1ull << sh) < d) ++sh;
    f.sh = sh;
    f.mul = (uint32_t)((((1ull << sh) - d) << 32) / d + 1);
    return f;
}
QEB_HD uint32_t fd_div(const FastDiv& f, uint32_t x)
{
#if defined(__CUDA_ARCH__)
    return (__umulhi(x, f.mul) + x) >> f.sh;
#else
    return (uint32_t)(((((uint64_t)x * f.mul) >> 32) + x) >> f.sh);
#endif
}

constexpr int QEB_MAXPASS = 8;

// One FFT of length N as a list of in-place radix passes over positions
// base + A*b (b < r), base = blk*L + a. DIF runs L = N, N/r0, ...; DIT the
// reverse. rev[pos] = transform index held at position pos after a DIF
// (= index expected at pos before a DIT).
struct FftPlan {
    int n;
    int npass;
    int r[QEB_MAXPASS];      // radix, DIF execution order
    int L[QEB_MAXPASS];      // sub-length of pass p (DIF order): L[0] = n
    FastDiv divA[QEB_MAXPASS];  // divide by A = L/r
    const cplx* tw;          // device: exp(+2 pi i j / n), j < n
    const int* rev;          // device: digit-reversal table, n entries
};

}  // namespace qeb
