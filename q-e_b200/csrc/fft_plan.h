// Host-side FFT plan construction: factorisation into the radices the
// kernels implement, exact twiddle table and digit-reversal table.
// Allowed lengths are those of the reference (factors 2, 3, 5 only:
// FFTXlib/src/fft_support.f90:61-120).
#pragma once
#include <math.h>
#include <vector>
#include "common.cuh"

namespace qeb {

struct HostPlan {
    int n = 0;
    std::vector<int> r;  // DIF order
    std::vector<int> L;
    std::vector<cplx> tw;  // exp(+2 pi i j/n)
    std::vector<int> rev;  // rev[pos] = k
};

// exp(2 pi i j / n) with octant reduction in long double
inline cplx exact_twiddle(int j, int n)
{
    long long num = j % n, den = n;
    long long q = (4 * num) / den;
    long long num4 = 4 * num - q * den;
    bool swap = false;
    if (2 * num4 > den) {
        num4 = den - num4;
        swap = true;
    }
    long double c, s;
    if (num4 == 0) {
        c = 1.0L;
        s = 0.0L;
    } else if (2 * num4 == den) {
        c = s = sqrtl(0.5L);
    } else {
        const long double ang = 3.14159265358979323846264338327950288L / 2.0L * (long double)num4 / (long double)den;
        c = cosl(ang);
        s = sinl(ang);
    }
    if (swap) {
        long double t = c;
        c = s;
        s = t;
    }
    for (long long i = 0; i < q; ++i) {
        long double t = c;
        c = -s;
        s = t;
    }
    return mk((double)c, (double)s);
}

// Radix choice: as few passes as possible with radices <= maxr from
// {10, 9, 8, 6, 5, 4, 3, 2}; among those the most balanced list (smallest sum
// of squares); large radices first (DIF order). fft_static.cuh instantiates
// the same lists for the common grid sizes.
inline bool factorize(int n, std::vector<int>& out, int maxr = 10)
{
    static const int cand[] = {10, 9, 8, 6, 5, 4, 3, 2};
    out.clear();
    if (n < 1) return false;
    if (n == 1) return true;
    const long BIG = 1L << 40;
    std::vector<long> best(n + 1, BIG);  // passes * 2^20 + sum of squares
    std::vector<int> pick(n + 1, 0);
    best[1] = 0;
    for (int m = 2; m <= n; ++m) {
        if (n % m) continue;
        for (int c : cand) {
            if (c > maxr || m % c) continue;
            if (best[m / c] >= BIG) continue;
            const long cost = best[m / c] + (1L << 20) + (long)c * c;
            if (cost < best[m]) {
                best[m] = cost;
                pick[m] = c;
            }
        }
    }
    if (best[n] >= BIG) return false;
    for (int m = n; m > 1; m /= pick[m]) out.push_back(pick[m]);
    for (size_t i = 0; i < out.size(); ++i)
        for (size_t k = i + 1; k < out.size(); ++k)
            if (out[k] > out[i]) {
                int t = out[i];
                out[i] = out[k];
                out[k] = t;
            }
    return (int)out.size() <= QEB_MAXPASS;
}

// Lengths with a compile-time plan in fft_static.cuh use its (large, in-register) radix list, so that
// the fast kernels and the generic fallback share one twiddle/digit-reversal table per length.
inline bool preferred_factors(int n, std::vector<int>& out)
{
    switch (n) {
    case 15: out = {15}; return true;
    case 25: out = {25}; return true;
    case 27: out = {27}; return true;
    case 64: out = {8, 8}; return true;
    case 125: out = {25, 5}; return true;
    case 135: out = {15, 9}; return true;
    case 144: out = {12, 12}; return true;
    case 200: out = {8, 5, 5}; return true;  // measured at AUSURF112 (z passes): 20.3 vs 20.7 us/band for {25, 8} (profiles/r2k_sweep.log)
    default: return false;
    }
}

// "8,5,5" -> {8,5,5} when the product is n and every radix has a butterfly; tuning aid for the QEB_[XYZ]PLAN
// environment overrides (an alternative radix list only takes the fast kernels if fft_static.cuh instantiates it)
inline bool parse_plan(const char* txt, int n, std::vector<int>& out)
{
    out.clear();
    if (!txt) return false;
    long prod = 1;
    for (const char* c = txt; *c;) {
        int r = 0;
        while (*c >= '0' && *c <= '9') r = r * 10 + (*c++ - '0');
        if (r < 2 || r > 27) return false;
        out.push_back(r);
        prod *= r;
        if (*c == ',') ++c;
        else if (*c) return false;
    }
    return prod == n && (int)out.size() <= QEB_MAXPASS;
}

inline bool make_host_plan(int n, HostPlan& hp, int maxr = 10, const char* override_txt = nullptr)
{
    hp.n = n;
    if (parse_plan(override_txt, n, hp.r)) {
        static const int ok[] = {2, 3, 4, 5, 6, 8, 9, 10, 12, 15, 16, 20, 25, 27};
        for (int r : hp.r) {
            bool f = false;
            for (int o : ok) f = f || (o == r);
            if (!f) return false;
        }
    } else if (!(maxr >= 10 && preferred_factors(n, hp.r)) && !factorize(n, hp.r, maxr))
        return false;
    if (hp.r.empty()) hp.r.push_back(1);  // n == 1: not used by the kernels
    hp.L.clear();
    int L = n;
    for (int r : hp.r) {
        hp.L.push_back(L);
        L /= r;
    }
    hp.tw.resize(n);
    for (int j = 0; j < n; ++j) hp.tw[j] = exact_twiddle(j, n);
    // rev: pos = sum d_p * A_p (A_p = L_p / r_p), k = d_0 + r_0 d_1 + r_0 r_1 d_2 ...
    hp.rev.assign(n, 0);
    for (int pos = 0; pos < n; ++pos) {
        int rem = pos, k = 0, mult = 1;
        for (size_t p = 0; p < hp.r.size(); ++p) {
            int A = hp.L[p] / hp.r[p];
            int d = rem / A;
            rem -= d * A;
            k += d * mult;
            mult *= hp.r[p];
        }
        hp.rev[pos] = k;
    }
    return true;
}

inline void fill_plan(const HostPlan& hp, const cplx* tw_dev, const int* rev_dev, FftPlan& pl)
{
    pl.n = hp.n;
    pl.npass = (int)hp.r.size();
    for (int p = 0; p < QEB_MAXPASS; ++p) {
        pl.r[p] = p < pl.npass ? hp.r[p] : 1;
        pl.L[p] = p < pl.npass ? hp.L[p] : 1;
        int A = pl.L[p] / pl.r[p];
        pl.divA[p] = make_fastdiv(A < 1 ? 1 : A);
    }
    pl.tw = tw_dev;
    pl.rev = rev_dev;
}

}  // namespace qeb
