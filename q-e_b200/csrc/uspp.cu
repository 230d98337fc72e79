// g2_kin and init_us_2 on the device: the per-k-point inputs of h_psi are produced where they are
// consumed, so a k-point change costs no 16*npwx*nkb-byte upload of vkb (3.08 GB at AUSURF112).
// Reference: PW/src/g2_kin.f90:10-62; upflib/init_us_2_acc.f90:10-172, upflib/ylmr2.f90:9-138 (CPU branch),
// upflib/beta_mod.f90:130-173 (interp_beta).
#include "uspp.h"

namespace qeb {

__global__ void k_g2_kin(int npw, const int* __restrict__ igk, const double* __restrict__ g, double xk1, double xk2,
                         double xk3, double tpiba2, double qcutz, double ecfixed, double q2sigma,
                         double* __restrict__ g2kin)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npw) return;
    const int iv = (igk ? igk[i] : i + 1) - 1;
    const double a = xk1 + g[3 * (size_t)iv], b = xk2 + g[3 * (size_t)iv + 1], c = xk3 + g[3 * (size_t)iv + 2];
    double v = (a * a + b * b + c * c) * tpiba2;
    if (qcutz > 0.0) v = v + qcutz * (1.0 + erf((v - ecfixed) / q2sigma));
    g2kin[i] = v;
}

cudaError_t launch_g2_kin(int npw, const int* igk, const double* g, double xk1, double xk2, double xk3, double tpiba2,
                          double qcutz, double ecfixed, double q2sigma, double* g2kin, cudaStream_t st)
{
    if (npw <= 0) return cudaSuccess;
    k_g2_kin<<<(npw + 255) / 256, 256, 0, st>>>(npw, igk, g, xk1, xk2, xk3, tpiba2, qcutz, ecfixed, q2sigma, g2kin);
    return cudaGetLastError();
}

constexpr int US_MAXL = 6;  // lmaxkb <= 6: (6+1)^2 = 49 harmonics per thread

__global__ void k_us2_ylm_vq(int npw, const int* __restrict__ igk, const double* __restrict__ g, double q1, double q2,
                             double q3, double tpiba, UsppTables u, double* __restrict__ ylm, double* __restrict__ vq)
{
    const int ig = blockIdx.x * blockDim.x + threadIdx.x;
    if (ig >= npw) return;
    const double pi = 3.14159265358979323846, fpi = 4.0 * pi, eps = 1.0e-9;
    const int iv = (igk ? igk[ig] : ig + 1) - 1;
    const double g1 = q1 + g[3 * (size_t)iv], g2 = q2 + g[3 * (size_t)iv + 1], g3 = q3 + g[3 * (size_t)iv + 2];
    const double gg = g1 * g1 + g2 * g2 + g3 * g3;
    const int lmax = u.lmaxkb, lmax2 = (lmax + 1) * (lmax + 1);
    double y[(US_MAXL + 1) * (US_MAXL + 1)];
    if (lmax == 0) {  // ylmr2.f90:55-58
        y[0] = sqrt(1.0 / fpi);
    } else {
        const double gmod = sqrt(gg);
        const double cost = gmod < eps ? 0.0 : g3 / gmod;
        const double sent = sqrt(fmax(0.0, 1.0 - cost * cost));
        // Q(l,m) = sqrt((l-m)!/(l+m)!) P(l,m) stored at lm = l^2 + 1 + 2m (1-based), ylmr2.f90:70-96
        y[0] = 1.0;
        y[1] = cost;
        y[3] = -sent / sqrt(2.0);
        for (int l = 2; l <= lmax; ++l) {
            for (int m = 0; m <= l - 2; ++m) {
                const int lm = l * l + 2 * m, lm1 = (l - 1) * (l - 1) + 2 * m, lm2 = (l - 2) * (l - 2) + 2 * m;
                y[lm] = cost * (2 * l - 1) / sqrt((double)(l * l - m * m)) * y[lm1] -
                        sqrt((double)((l - 1) * (l - 1) - m * m)) / sqrt((double)(l * l - m * m)) * y[lm2];
            }
            const int lm = l * l + 2 * l, lm1 = l * l + 2 * (l - 1), lm2 = (l - 1) * (l - 1) + 2 * (l - 1);
            y[lm1] = cost * sqrt((double)(2 * l - 1)) * y[lm2];
            y[lm] = -sqrt((double)(2 * l - 1)) / sqrt((double)(2 * l)) * sent * y[lm2];
        }
        double phi;  // ylmr2.f90:101-107
        if (g1 > eps)
            phi = atan(g2 / g1);
        else if (g1 < -eps)
            phi = atan(g2 / g1) + pi;
        else
            phi = copysign(pi / 2.0, g2);
        y[0] = y[0] / sqrt(fpi);
        int lm = 0;  // 0-based index of the last harmonic written
        for (int l = 1; l <= lmax; ++l) {
            const double c = sqrt((double)(2 * l + 1) / fpi);
            lm += 1;
            y[lm] = c * y[lm];
            for (int m = 1; m <= l; ++m) {
                lm += 2;
                const double q = y[lm];  // Q(l,m) sits at the "m < 0" slot
                y[lm - 1] = c * sqrt(2.0) * q * cos(m * phi);
                y[lm] = c * sqrt(2.0) * q * sin(m * phi);
            }
        }
    }
    for (int k = 0; k < lmax2; ++k) ylm[(size_t)k * npw + ig] = y[k];
    // interp_beta (beta_mod.f90:147-170) at |q+G| in atomic units
    const double qgr = sqrt(gg) * tpiba;
    const double t = qgr / u.dq;
    const int it = (int)t;
    const double px = t - (double)it, ux = 1.0 - px, vx = 2.0 - px, wx = 3.0 - px;
    const int i0 = it;  // 0-based (reference: INT(qgr/dq) + 1)
    const bool inside = (i0 + 4) <= u.nqx;
    for (int nt = 0; nt < u.nsp; ++nt)
        for (int nb = 0; nb < u.nbetam; ++nb) {
            const double* tb = u.tab_beta + ((size_t)nt * u.nbetam + nb) * u.nqx;
            double v = 0.0;
            if (inside)
                v = tb[i0] * ux * vx * wx / 6.0 + tb[i0 + 1] * px * vx * wx / 2.0 - tb[i0 + 2] * px * ux * wx / 2.0 +
                    tb[i0 + 3] * px * ux * vx / 6.0;
            vq[((size_t)nt * u.nbetam + nb) * npw + ig] = v;
        }
}

cudaError_t launch_us2_ylm_vq(int npw, const int* igk, const double* g, double q1, double q2, double q3, double tpiba,
                              const UsppTables& u, double* ylm, double* vq, cudaStream_t st)
{
    if (npw <= 0) return cudaSuccess;
    if (u.lmaxkb < 0 || u.lmaxkb > US_MAXL) return cudaErrorInvalidValue;
    k_us2_ylm_vq<<<(npw + 127) / 128, 128, 0, st>>>(npw, igk, g, q1, q2, q3, tpiba, u, ylm, vq);
    return cudaGetLastError();
}

// grid (ig tiles, atoms): every thread owns one plane wave of one atom and writes that atom's nh columns of vkb
// (coalesced along ig). Rows npw..npwx-1 are zeroed (vkb_(:,:) = 0, init_us_2_acc.f90:89-91).
__global__ void k_us2_vkb(int npw, int npwx, const int* __restrict__ igk, const int* __restrict__ mill,
                          const int* __restrict__ ityp, const int* __restrict__ jkb0, const cplx* __restrict__ qphase,
                          int nr1, int nr2, int nr3, const cplx* __restrict__ eigts1, const cplx* __restrict__ eigts2,
                          const cplx* __restrict__ eigts3, UsppTables u, const double* __restrict__ ylm,
                          const double* __restrict__ vq, cplx* __restrict__ vkb)
{
    const int ig = blockIdx.x * blockDim.x + threadIdx.x;
    const int na = blockIdx.y;
    if (ig >= npwx) return;
    const int nt = ityp[na] - 1;
    const int nh = u.nh[nt];
    cplx* out = vkb + (size_t)jkb0[na] * npwx + ig;
    if (ig >= npw) {
        for (int ih = 0; ih < nh; ++ih) out[(size_t)ih * npwx] = mk(0.0, 0.0);
        return;
    }
    const int iv = (igk ? igk[ig] : ig + 1) - 1;
    const int m1 = mill[3 * (size_t)iv], m2 = mill[3 * (size_t)iv + 1], m3 = mill[3 * (size_t)iv + 2];
    const cplx e1 = eigts1[(size_t)na * (2 * nr1 + 1) + m1 + nr1];
    const cplx e2 = eigts2[(size_t)na * (2 * nr2 + 1) + m2 + nr2];
    const cplx e3 = eigts3[(size_t)na * (2 * nr3 + 1) + m3 + nr3];
    const cplx sk = cmul(cmul(cmul(e1, e2), e3), qphase[na]);  // init_us_2_acc.f90:140-143
    for (int ih = 0; ih < nh; ++ih) {
        const int l = u.nhtol[nt * u.nhm + ih];
        const int lm = u.nhtolm[nt * u.nhm + ih] - 1;
        const int nb = u.indv[nt * u.nhm + ih] - 1;
        const double v1 = ylm[(size_t)lm * npw + ig] * vq[((size_t)nt * u.nbetam + nb) * npw + ig];  // :119
        const cplx w = mk(v1 * sk.x, v1 * sk.y);
        cplx r;  // times (-i)^l (:150)
        switch (l & 3) {
        case 0: r = w; break;
        case 1: r = mk(w.y, -w.x); break;
        case 2: r = mk(-w.x, -w.y); break;
        default: r = mk(-w.y, w.x); break;
        }
        out[(size_t)ih * npwx] = r;
    }
}

cudaError_t launch_us2_vkb(int npw, int npwx, const int* igk, const int* mill, int nat, const int* ityp,
                           const int* jkb0, const cplx* qphase, int nr1, int nr2, int nr3, const cplx* eigts1,
                           const cplx* eigts2, const cplx* eigts3, const UsppTables& u, const double* ylm,
                           const double* vq, cplx* vkb, cudaStream_t st)
{
    if (npwx <= 0 || nat <= 0) return cudaSuccess;
    dim3 grid((npwx + 255) / 256, nat);
    k_us2_vkb<<<grid, 256, 0, st>>>(npw, npwx, igk, mill, ityp, jkb0, qphase, nr1, nr2, nr3, eigts1, eigts2, eigts3, u,
                                    ylm, vq, vkb);
    return cudaGetLastError();
}

}  // namespace qeb
