/*
 * ORACLE / CPU BASELINE (test infrastructure, not product code).
 *
 * Restated CPU driver of the reference's V_loc application that executes
 * every 1-D transform with the reference's OWN C FFT engine (the internal
 * FFTW-2 codelets of /root/reference/FFTXlib/src/fftw_dp.c, reached through
 * the fft_stick.c wrappers create_plan_1d / fftw_inplace_drv_1d).  This file
 * is compiled TOGETHER with those reference sources, where they lie, into
 * oracle/_ref/libqe_ref.so (see oracle/Makefile); no reference source is
 * copied into this repository.
 *
 * Restates:
 *   cfft3ds          /root/reference/FFTXlib/src/fft_scalar.FFTW.f90:551-721
 *   fftx_c2psi_k     /root/reference/FFTXlib/src/fft_helper_subroutines.f90:447-511
 *   fftx_psi2c_k     /root/reference/FFTXlib/src/fft_helper_subroutines.f90:751-795
 *   fftx_c2psi_gamma /root/reference/FFTXlib/src/fft_helper_subroutines.f90:317-409
 *   fftx_psi2c_gamma /root/reference/FFTXlib/src/fft_helper_subroutines.f90:644-715
 *   vloc_psi_k_acc   /root/reference/PW/src/vloc_psi_acc.f90:148-266
 *   vloc_psi_gamma_acc /root/reference/PW/src/vloc_psi_acc.f90:9-145
 * Bands (k) / band pairs (Gamma) are spread over OpenMP threads, mirroring
 * the reference's band-group parallelism (-nb).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library.
 */
#include <stdlib.h>
#include <string.h>
#include <complex.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "fftw_dp.h"

/* reference wrappers, fft_stick.c:16-146 */
int create_plan_1d(fftw_plan *p, int *n, int *idir);
int destroy_plan_1d(fftw_plan *p);
int fftw_inplace_drv_1d(fftw_plan *p, int *nfft, FFTW_COMPLEX *a, int *inca, int *idist);

typedef double _Complex cplx;

typedef struct {
    int nx, ny, nz;
    fftw_plan fw[3], bw[3];
} plans3;

static void plans_make(plans3 *P, int nx, int ny, int nz)
{
    int d[3] = {nx, ny, nz}, i, idir;
    P->nx = nx; P->ny = ny; P->nz = nz;
    for (i = 0; i < 3; ++i) {
        idir = -1; create_plan_1d(&P->fw[i], &d[i], &idir);
        idir = 1;  create_plan_1d(&P->bw[i], &d[i], &idir);
    }
}
static void plans_free(plans3 *P)
{
    for (int i = 0; i < 3; ++i) { destroy_plan_1d(&P->fw[i]); destroy_plan_1d(&P->bw[i]); }
}

/* One "reduced" 3-D transform, reference order (fft_scalar.FFTW.f90:603-684).
 * f is the x-fastest box (ldx,ldy,ldz); do_fft_z has ldx*ldy entries,
 * do_fft_y has nx entries. isign>0: G->R, exp(+i), unscaled;
 * isign<0: R->G, exp(-i), then scaled by 1/(nx*ny*nz). */
static void cfft3ds_one(plans3 *P, cplx *f, int ldx, int ldy, int ldz, int isign,
                        const int *do_fft_z, const int *do_fft_y)
{
    int nx = P->nx, ny = P->ny, nz = P->nz;
    int i, j, m, incx1, incx2;
    (void)ldz;
    if (isign > 0) {
        incx1 = ldx * ldy; incx2 = 1; m = 1;
        for (i = 0; i < nx; ++i)
            for (j = 0; j < ny; ++j) {
                int ii = i + ldx * j;
                if (do_fft_z[ii] > 0)
                    fftw_inplace_drv_1d(&P->bw[2], &m, (FFTW_COMPLEX *)(f + ii), &incx1, &incx2);
            }
        incx1 = ldx; incx2 = ldx * ldy; m = nz;
        for (i = 0; i < nx; ++i)
            if (do_fft_y[i] == 1)
                fftw_inplace_drv_1d(&P->bw[1], &m, (FFTW_COMPLEX *)(f + i), &incx1, &incx2);
        incx1 = 1; incx2 = ldx; m = ldy * nz;
        fftw_inplace_drv_1d(&P->bw[0], &m, (FFTW_COMPLEX *)f, &incx1, &incx2);
    } else {
        incx1 = 1; incx2 = ldx; m = ldy * nz;
        fftw_inplace_drv_1d(&P->fw[0], &m, (FFTW_COMPLEX *)f, &incx1, &incx2);
        incx1 = ldx; incx2 = ldx * ldy; m = nz;
        for (i = 0; i < nx; ++i)
            if (do_fft_y[i] == 1)
                fftw_inplace_drv_1d(&P->fw[1], &m, (FFTW_COMPLEX *)(f + i), &incx1, &incx2);
        incx1 = ldx * ny; incx2 = 1; m = 1;
        for (i = 0; i < nx; ++i)
            for (j = 0; j < ny; ++j) {
                int ii = i + ldx * j;
                if (do_fft_z[ii] > 0)
                    fftw_inplace_drv_1d(&P->fw[2], &m, (FFTW_COMPLEX *)(f + ii), &incx1, &incx2);
            }
        double s = 1.0 / ((double)nx * ny * nz);
        long tot = (long)ldx * ldy * nz;
        for (long t = 0; t < tot; ++t) f[t] *= s;
    }
}

/* public: in-place reduced 3-D FFT of `howmany` consecutive boxes */
int qeref_cfft3ds(cplx *f, int nx, int ny, int nz, int ldx, int ldy, int ldz, int howmany,
                  int isign, const int *do_fft_z, const int *do_fft_y)
{
    plans3 P;
    if (ny != ldy || howmany < 1) return 1;
    plans_make(&P, nx, ny, nz);
    for (int h = 0; h < howmany; ++h)
        cfft3ds_one(&P, f + (long)h * ldx * ldy * ldz, ldx, ldy, ldz, isign, do_fft_z, do_fft_y);
    plans_free(&P);
    return 0;
}

/* hpsi(1:n,b) += [fwfft(v * invfft(scatter psi(:,b)))](nl(igk)) ; nl/igk 1-based */
int qeref_vloc_psi_k(int lda, int n, int m, const cplx *psi, const double *v, cplx *hpsi,
                     int nr1, int nr2, int nr3, const int *nl, const int *igk,
                     const int *do_fft_z, const int *do_fft_y, int nthreads)
{
    long nnr = (long)nr1 * nr2 * nr3;
    plans3 P;
    plans_make(&P, nr1, nr2, nr3);
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel
    {
        cplx *box = (cplx *)malloc(sizeof(cplx) * nnr);
#pragma omp for schedule(dynamic, 1)
        for (int b = 0; b < m; ++b) {
            const cplx *c = psi + (long)b * lda;
            cplx *h = hpsi + (long)b * lda;
            memset(box, 0, sizeof(cplx) * nnr);
            for (int j = 0; j < n; ++j) box[nl[igk[j] - 1] - 1] = c[j];
            cfft3ds_one(&P, box, nr1, nr2, nr3, +1, do_fft_z, do_fft_y);
            for (long r = 0; r < nnr; ++r) box[r] *= v[r];
            cfft3ds_one(&P, box, nr1, nr2, nr3, -1, do_fft_z, do_fft_y);
            for (int j = 0; j < n; ++j) h[j] += box[nl[igk[j] - 1] - 1];
        }
        free(box);
    }
    plans_free(&P);
    return 0;
}

/* Gamma: two bands per complex box (batched branch, vloc_psi_acc.f90:58-101) */
int qeref_vloc_psi_gamma(int lda, int n, int m, const cplx *psi, const double *v, cplx *hpsi,
                         int nr1, int nr2, int nr3, const int *nl, const int *nlm,
                         const int *do_fft_z, const int *do_fft_y, int nthreads)
{
    long nnr = (long)nr1 * nr2 * nr3;
    int npair = (m + 1) / 2;
    plans3 P;
    plans_make(&P, nr1, nr2, nr3);
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel
    {
        cplx *box = (cplx *)malloc(sizeof(cplx) * nnr);
#pragma omp for schedule(dynamic, 1)
        for (int p = 0; p < npair; ++p) {
            int b1 = 2 * p, b2 = 2 * p + 1;
            int pair = (b2 < m);
            const cplx *c1 = psi + (long)b1 * lda;
            const cplx *c2 = pair ? psi + (long)b2 * lda : NULL;
            memset(box, 0, sizeof(cplx) * nnr);
            if (pair) {
                for (int j = 0; j < n; ++j) {
                    box[nl[j] - 1] = c1[j] + I * c2[j];
                    box[nlm[j] - 1] = conj(c1[j] - I * c2[j]);
                }
            } else {
                for (int j = 0; j < n; ++j) {
                    box[nl[j] - 1] = c1[j];
                    box[nlm[j] - 1] = conj(c1[j]);
                }
            }
            cfft3ds_one(&P, box, nr1, nr2, nr3, +1, do_fft_z, do_fft_y);
            for (long r = 0; r < nnr; ++r) box[r] *= v[r];
            cfft3ds_one(&P, box, nr1, nr2, nr3, -1, do_fft_z, do_fft_y);
            cplx *h1 = hpsi + (long)b1 * lda;
            if (pair) {
                cplx *h2 = hpsi + (long)b2 * lda;
                for (int j = 0; j < n; ++j) {
                    cplx a = box[nl[j] - 1], bm = box[nlm[j] - 1];
                    cplx fp = (a + bm) * 0.5, fm = (a - bm) * 0.5;
                    h1[j] += creal(fp) + I * cimag(fm);
                    h2[j] += cimag(fp) - I * creal(fm);
                }
            } else {
                for (int j = 0; j < n; ++j) h1[j] += box[nl[j] - 1];
            }
        }
        free(box);
    }
    plans_free(&P);
    return 0;
}
