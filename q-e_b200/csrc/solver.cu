// Device-resident block-Davidson solver: qeb_cegterg (KS_Solvers/Davidson/cegterg.f90:20-700, k-points) and
// qeb_regterg (regterg.f90:20-640, Gamma) for one band group and serial diagonalisation.
//
// The reference allocates psi, hpsi, spsi (npwx x nvecx) and the reduced matrices hc, sc, vc per call and reaches
// H|psi>, S|psi> and the preconditioner through EXTERNAL pointers (cegterg.f90:110-118). Here the three work arrays
// and the reduced matrices live in HBM for the whole solve, the operators are the context's own (qeb_h_psi, qeb_s_psi,
// qeb_g_psi), the reduced matrices come from the DMMA projector kernels (qeb_dav_gemm_cn / _nn) and the nbase x nbase
// generalized eigenproblem (`diaghg`, LAPACK in the reference, cuSOLVER in its GPU build) is cusolverDnZhegvd /
// cusolverDnDsygvd, resolved at run time with dlopen like NCCL. Per iteration only the nvec eigenvalue estimates cross
// to the host (the convergence bookkeeping of cegterg.f90:556-575 is host code); evc crosses once in and once out.
#include <cusolverDn.h>
#include <dlfcn.h>
#include "context.h"

namespace {

struct SolverApi {
    bool ok = false;
    const char* why = "";
    cusolverStatus_t (*Create)(cusolverDnHandle_t*) = nullptr;
    cusolverStatus_t (*Destroy)(cusolverDnHandle_t) = nullptr;
    cusolverStatus_t (*SetStream)(cusolverDnHandle_t, cudaStream_t) = nullptr;
    cusolverStatus_t (*ZhegvdBuf)(cusolverDnHandle_t, cusolverEigType_t, cusolverEigMode_t, cublasFillMode_t, int,
                                  const cuDoubleComplex*, int, const cuDoubleComplex*, int, const double*, int*) = nullptr;
    cusolverStatus_t (*Zhegvd)(cusolverDnHandle_t, cusolverEigType_t, cusolverEigMode_t, cublasFillMode_t, int, cuDoubleComplex*,
                               int, cuDoubleComplex*, int, double*, cuDoubleComplex*, int, int*) = nullptr;
    cusolverStatus_t (*DsygvdBuf)(cusolverDnHandle_t, cusolverEigType_t, cusolverEigMode_t, cublasFillMode_t, int, const double*,
                                  int, const double*, int, const double*, int*) = nullptr;
    cusolverStatus_t (*Dsygvd)(cusolverDnHandle_t, cusolverEigType_t, cusolverEigMode_t, cublasFillMode_t, int, double*, int,
                               double*, int, double*, double*, int, int*) = nullptr;
};

const SolverApi& solver_api()
{
    static SolverApi A;
    static bool tried = false;
    if (tried) return A;
    tried = true;
    void* h = nullptr;
    if (const char* p = getenv("QEB_CUSOLVER_LIB")) h = dlopen(p, RTLD_NOW | RTLD_GLOBAL);
    const char* names[] = {"libcusolver.so.11", "libcusolver.so", "/usr/local/cuda/lib64/libcusolver.so.11",
                           "/usr/local/cuda/lib64/libcusolver.so"};
    for (const char* n : names)
        if (!h) h = dlopen(n, RTLD_NOW | RTLD_NOLOAD);
    for (const char* n : names)
        if (!h) h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (!h) {
        A.why = "libcusolver not found (set QEB_CUSOLVER_LIB)";
        return A;
    }
#define QEB_SYM(field, name)                            \
    *(void**)(&A.field) = dlsym(h, name);               \
    if (!A.field) {                                     \
        A.why = "symbol " name " missing in libcusolver"; \
        return A;                                       \
    }
    QEB_SYM(Create, "cusolverDnCreate")
    QEB_SYM(Destroy, "cusolverDnDestroy")
    QEB_SYM(SetStream, "cusolverDnSetStream")
    QEB_SYM(ZhegvdBuf, "cusolverDnZhegvd_bufferSize")
    QEB_SYM(Zhegvd, "cusolverDnZhegvd")
    QEB_SYM(DsygvdBuf, "cusolverDnDsygvd_bufferSize")
    QEB_SYM(Dsygvd, "cusolverDnDsygvd")
#undef QEB_SYM
    A.ok = true;
    return A;
}

// ---- small kernels on the reduced matrices (column-major, leading dimension ld; CPLX: complex elements) ---------
// hc(n,m) = conj(hc(m,n)) for m >= max(n+1, lo), diagonal n >= lo strictly real (cegterg.f90:241-254, 520-533)
template <bool CPLX>
__global__ void k_dav_symmetrise(double* a, int ld, int lo, int hi)
{
    const int m = blockIdx.x * blockDim.x + threadIdx.x;  // row of the master (lower) element
    const int n = blockIdx.y;                             // its column
    if (m >= hi || n >= hi) return;
    constexpr int E = CPLX ? 2 : 1;
    if (m == n) {
        if (CPLX && n >= lo) a[((size_t)n * ld + n) * E + 1] = 0.0;
        return;
    }
    if (m < n || m < lo) return;
    const double* src = a + ((size_t)n * ld + m) * E;  // (m, n)
    double* dst = a + ((size_t)m * ld + n) * E;        // (n, m)
    dst[0] = src[0];
    if (CPLX) dst[1] = -src[1];
}

// dst(:, k) = src(:, idx[k]) for k < m (rows < n)
template <bool CPLX>
__global__ void k_dav_gather_cols(const double* src, double* dst, int ld, int n, int m, const int* idx)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int k = blockIdx.y;
    if (i >= n || k >= m) return;
    constexpr int E = CPLX ? 2 : 1;
    const double* s = src + ((size_t)idx[k] * ld + i) * E;
    double* d = dst + ((size_t)k * ld + i) * E;
    d[0] = s[0];
    if (CPLX) d[1] = s[1];
}

// refresh of the reduced matrices (cegterg.f90:640-668): hc = diag(e), sc = vc = 1 on the leading n x n block
template <bool CPLX>
__global__ void k_dav_reset(double* hc, double* sc, double* vc, int ld, int n, const double* e)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= n || j >= n) return;
    constexpr int E = CPLX ? 2 : 1;
    const size_t o = ((size_t)j * ld + i) * E;
    const double d = i == j ? 1.0 : 0.0;
    hc[o] = i == j ? e[i] : 0.0;
    sc[o] = d;
    vc[o] = d;
    if (CPLX) hc[o + 1] = sc[o + 1] = vc[o + 1] = 0.0;
}

// psi(:, k) *= f(s[k]) on rows < n.  MODE 0: -s[k] (cegterg.f90:356-375); MODE 1: 1/sqrt(s[k]) (:436-458)
template <int MODE>
__global__ void k_dav_scale_cols(cplx* psi, int lda, int n, int m, const double* s)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int k = blockIdx.y;
    if (i >= n || k >= m) return;
    const double f = MODE == 0 ? -s[k] : 1.0 / sqrt(s[k]);
    cplx* p = psi + (size_t)k * lda + i;
    const cplx v = *p;
    *p = mk(v.x * f, v.y * f);
}

// v = identity on the leading n x n block of a zeroed matrix
template <bool CPLX>
__global__ void k_dav_identity(double* v, int ld, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) v[((size_t)i * ld + i) * (CPLX ? 2 : 1)] = 1.0;
}

template <bool CPLX>
__global__ void k_dav_diag(const double* hc, int ld, int n, double* e)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) e[i] = hc[((size_t)i * ld + i) * (CPLX ? 2 : 1)];
}

struct DavBufs {  // the solver's device storage, kept in the context between calls
    DevBuf psi, hpsi, spsi, hc, sc, vc, wa, wb, work, ew, evc_stage, idx, info;
    cusolverDnHandle_t handle = nullptr;
};

}  // namespace

struct qeb_solver_state {
    DavBufs B;
};

static int dav_fail(qeb_ctx* c, int code, const std::string& msg)
{
    c->err = msg;
    return code;
}

// diaghg(n, m, h, s, ldh, e, v): lowest m eigenpairs of the leading n x n blocks, eigenvectors in v(:, 1:m), eigenvalues to
// the host array e_h (LAXlib cdiaghg / rdiaghg; the reference's GPU build calls cusolverDnZhegvdx there)
static int dav_diaghg(qeb_ctx* c, DavBufs& B, int n, int m, int ld, bool cplx_, double* e_h)
{
    const SolverApi& A = solver_api();
    const size_t el = cplx_ ? 16 : 8;
    QEB_CUDA(c, cudaMemcpy2DAsync(B.wa.p, (size_t)ld * el, B.hc.p, (size_t)ld * el, (size_t)n * el, n, cudaMemcpyDeviceToDevice, c->stream));
    QEB_CUDA(c, cudaMemcpy2DAsync(B.wb.p, (size_t)ld * el, B.sc.p, (size_t)ld * el, (size_t)n * el, n, cudaMemcpyDeviceToDevice, c->stream));
    int lwork = 0;
    cusolverStatus_t st;
    if (cplx_)
        st = A.ZhegvdBuf(B.handle, CUSOLVER_EIG_TYPE_1, CUSOLVER_EIG_MODE_VECTOR, CUBLAS_FILL_MODE_LOWER, n, (cuDoubleComplex*)B.wa.p,
                         ld, (cuDoubleComplex*)B.wb.p, ld, (double*)B.ew.p, &lwork);
    else
        st = A.DsygvdBuf(B.handle, CUSOLVER_EIG_TYPE_1, CUSOLVER_EIG_MODE_VECTOR, CUBLAS_FILL_MODE_LOWER, n, (double*)B.wa.p, ld,
                         (double*)B.wb.p, ld, (double*)B.ew.p, &lwork);
    if (st != CUSOLVER_STATUS_SUCCESS) return dav_fail(c, 120, "diaghg: cusolver bufferSize failed (status " + std::to_string((int)st) + ")");
    QEB_CUDA(c, B.work.ensure((size_t)lwork * el));
    if (cplx_)
        st = A.Zhegvd(B.handle, CUSOLVER_EIG_TYPE_1, CUSOLVER_EIG_MODE_VECTOR, CUBLAS_FILL_MODE_LOWER, n, (cuDoubleComplex*)B.wa.p, ld,
                      (cuDoubleComplex*)B.wb.p, ld, (double*)B.ew.p, (cuDoubleComplex*)B.work.p, lwork, (int*)B.info.p);
    else
        st = A.Dsygvd(B.handle, CUSOLVER_EIG_TYPE_1, CUSOLVER_EIG_MODE_VECTOR, CUBLAS_FILL_MODE_LOWER, n, (double*)B.wa.p, ld,
                      (double*)B.wb.p, ld, (double*)B.ew.p, (double*)B.work.p, lwork, (int*)B.info.p);
    if (st != CUSOLVER_STATUS_SUCCESS) return dav_fail(c, 121, "diaghg: cusolverDn*gvd failed (status " + std::to_string((int)st) + ")");
    // v(:, 1:m) <- the m lowest eigenvectors; e <- the m lowest eigenvalues
    QEB_CUDA(c, cudaMemcpy2DAsync(B.vc.p, (size_t)ld * el, B.wa.p, (size_t)ld * el, (size_t)n * el, m, cudaMemcpyDeviceToDevice, c->stream));
    int info = 0;
    QEB_CUDA(c, cudaMemcpyAsync(e_h, B.ew.p, (size_t)m * 8, cudaMemcpyDeviceToHost, c->stream));
    QEB_CUDA(c, cudaMemcpyAsync(&info, B.info.p, 4, cudaMemcpyDeviceToHost, c->stream));
    QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    if (info != 0)  // the reference: errore('cdiaghg', 'S matrix not positive definite' / 'eigenvectors failed to converge')
        return dav_fail(c, 122, "diaghg: cusolver info = " + std::to_string(info) + (info > n ? " (S matrix not positive definite)" : ""));
    return 0;
}

static int dav_solve(qeb_ctx* c, int npw, int npwx, int nvec, int nvecx, int npol, void* evc, double ethr, int uspp, double* e,
                     const int* btype, int* notcnv_out, int lrot, int* dav_iter_out, int* nhpsi_out, const char* who)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (!c->have_k) QEB_FAIL(c, 60, std::string(who) + ": qeb_set_kpoint has not been called");
    if (npol != 1 && npol != 2) QEB_FAIL(c, 110, std::string(who) + ": npol must be 1 or 2");
    if (npol == 2 && c->gamma) QEB_FAIL(c, 110, std::string(who) + ": npol = 2 on a gamma_only grid");
    if (c->dist && c->DH.nproc > 1)
        QEB_FAIL(c, 111, std::string(who) + ": slab-distributed grids are not built (mp_sum of the reduced matrices over intra_bgrp_comm)");
    if (npw != c->npw || npwx < npw || nvec < 1 || !evc || !e || !btype)
        QEB_FAIL(c, 112, std::string(who) + ": bad arguments (npw must be the npw of qeb_set_kpoint)");
    if (nvec > nvecx / 2) QEB_FAIL(c, 113, std::string(who) + ": nvecx is too small");  // cegterg.f90:125
    if (npol == 1 ? c->gpsi_n < npw : c->gpsi_n != 2 * npwx)
        QEB_FAIL(c, 114, std::string(who) + ": no preconditioner diagonals (qeb_set_g_psi_diag; npol = 2: n = npwx*npol)");
    // cegterg.f90:131-142: spinors are (npwx*npol) long and every product runs over the pad rows too
    const int kdim = npol == 1 ? npw : npwx * npol;
    const int kdmx = npwx * npol;
    auto apply_h = [&](const cplx* x, cplx* hx, int m) {
        return npol == 1 ? qeb_h_psi(c, npwx, npw, m, x, hx) : qeb_h_psi_nc(c, npwx, npw, m, x, hx);
    };
    auto apply_s = [&](const cplx* x, cplx* sx, int m) {
        return npol == 1 ? qeb_s_psi(c, npwx, npw, m, x, sx) : qeb_s_psi_nc(c, npwx, npw, m, x, sx);
    };
    const SolverApi& A = solver_api();
    if (!A.ok) QEB_FAIL(c, 115, std::string(who) + ": " + A.why);
    if (!c->solver) c->solver = new qeb_solver_state();
    DavBufs& B = c->solver->B;
    if (!B.handle && A.Create(&B.handle) != CUSOLVER_STATUS_SUCCESS) QEB_FAIL(c, 116, std::string(who) + ": cusolverDnCreate failed");
    A.SetStream(B.handle, c->stream);
    const bool cp = !c->gamma;
    const size_t el = cp ? 16 : 8;
    const size_t col = (size_t)kdmx * 16;
    QEB_CUDA(c, B.psi.ensure(col * nvecx));
    QEB_CUDA(c, B.hpsi.ensure(col * nvecx));
    if (uspp) QEB_CUDA(c, B.spsi.ensure(col * nvecx));
    const size_t red = (size_t)nvecx * nvecx * el;
    QEB_CUDA(c, B.hc.ensure(red));
    QEB_CUDA(c, B.sc.ensure(red));
    QEB_CUDA(c, B.vc.ensure(red));
    QEB_CUDA(c, B.wa.ensure(red));
    QEB_CUDA(c, B.wb.ensure(red));
    QEB_CUDA(c, B.ew.ensure((size_t)nvecx * 8));
    QEB_CUDA(c, B.idx.ensure((size_t)nvecx * 4));
    QEB_CUDA(c, B.info.ensure(16));
    cplx* psi = (cplx*)B.psi.p;
    cplx* hpsi = (cplx*)B.hpsi.p;
    cplx* spsi = uspp ? (cplx*)B.spsi.p : nullptr;
    cplx* sb = uspp ? spsi : psi;
    double* ew_d = (double*)B.ew.p;
    const bool evc_dev = is_device_ptr(evc);
    cplx* evc_d = (cplx*)evc;
    if (!evc_dev) {
        QEB_CUDA(c, B.evc_stage.ensure(col * nvec));
        evc_d = (cplx*)B.evc_stage.p;
        QEB_CUDA(c, cudaMemcpyAsync(evc_d, evc, col * nvec, cudaMemcpyHostToDevice, c->stream));
    }
    const bool was_async = c->async;
    c->async = true;  // the operators below queue on the context's stream; the solver synchronises where it reads back
    struct Restore {
        qeb_ctx* c;
        bool a;
        ~Restore() { c->async = a; }
    } restore{c, was_async};
    auto colp = [&](cplx* base, int k) { return base + (size_t)k * kdmx; };
    auto red_at = [&](DevBuf& b, int row, int colm) { return (char*)b.p + ((size_t)colm * nvecx + row) * el; };
    const dim3 tb(256);
    int rc;
    std::vector<double> ew(nvecx, 0.0);
    std::vector<char> conv(nvec, 0);
    std::vector<int> idx(nvecx);
    const double empty_ethr = std::max(ethr * 5.0, 1e-5);  // cegterg.f90:128
    int notcnv = nvec, nbase = nvec, nhpsi = 0, dav_iter = 0;
    QEB_CUDA(c, cudaMemsetAsync(B.hc.p, 0, red, c->stream));
    QEB_CUDA(c, cudaMemsetAsync(B.sc.p, 0, red, c->stream));
    QEB_CUDA(c, cudaMemsetAsync(B.vc.p, 0, red, c->stream));
    QEB_CUDA(c, cudaMemcpyAsync(psi, evc_d, col * nvec, cudaMemcpyDeviceToDevice, c->stream));  // :190
    if (!cp && c->gstart == 2) QEB_CUDA(c, launch_zero_im_g0(psi, npwx, nvec, c->stream));       // regterg.f90:174
    if ((rc = apply_h(psi, hpsi, nvec))) return rc;                                          // :192
    nhpsi += nvec;
    if (uspp && (rc = apply_s(psi, spsi, nvec))) return rc;                                 // :196
    // hc = psi^H hpsi, sc = psi^H spsi (:209-229)
    if ((rc = qeb_dav_gemm_cn(c, kdim, psi, kdmx, nbase, hpsi, kdmx, nbase, B.hc.p, nvecx))) return rc;
    if ((rc = qeb_dav_gemm_cn(c, kdim, psi, kdmx, nbase, sb, kdmx, nbase, B.sc.p, nvecx))) return rc;
    auto symmetrise = [&](int lo, int hi) {
        const dim3 g((hi + 255) / 256, hi);
        if (cp) {
            k_dav_symmetrise<true><<<g, tb, 0, c->stream>>>((double*)B.hc.p, nvecx, lo, hi);
            k_dav_symmetrise<true><<<g, tb, 0, c->stream>>>((double*)B.sc.p, nvecx, lo, hi);
        } else {
            k_dav_symmetrise<false><<<g, tb, 0, c->stream>>>((double*)B.hc.p, nvecx, lo, hi);
            k_dav_symmetrise<false><<<g, tb, 0, c->stream>>>((double*)B.sc.p, nvecx, lo, hi);
        }
        return cudaGetLastError();
    };
    // the lower triangle is the master copy and the diagonal is made strictly real, as in the reference (:241-254)
    QEB_CUDA(c, symmetrise(0, nbase));
    if (lrot) {  // :257-267: vc = 1, e = diag(hc)
        if (cp) {
            k_dav_diag<true><<<(nbase + 255) / 256, tb, 0, c->stream>>>((double*)B.hc.p, nvecx, nbase, ew_d);
            k_dav_identity<true><<<(nbase + 255) / 256, tb, 0, c->stream>>>((double*)B.vc.p, nvecx, nbase);
        } else {
            k_dav_diag<false><<<(nbase + 255) / 256, tb, 0, c->stream>>>((double*)B.hc.p, nvecx, nbase, ew_d);
            k_dav_identity<false><<<(nbase + 255) / 256, tb, 0, c->stream>>>((double*)B.vc.p, nvecx, nbase);
        }
        QEB_CUDA(c, cudaGetLastError());
        QEB_CUDA(c, cudaMemcpyAsync(e, ew_d, (size_t)nvec * 8, cudaMemcpyDeviceToHost, c->stream));
        QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    } else {
        if ((rc = dav_diaghg(c, B, nbase, nvec, nvecx, cp, ew.data()))) return rc;  // :271-285
        for (int i = 0; i < nvec; ++i) e[i] = ew[i];
    }
    for (int kter = 1; kter <= 20; ++kter) {  // maxter = 20, cegterg.f90:71
        dav_iter = kter;
        // coefficients of the unconverged roots first (:305-334)
        int np = 0;
        bool moved = false;
        for (int n = 0; n < nvec; ++n)
            if (!conv[n]) {
                if (np != n) moved = true;
                idx[np] = n;
                ew[nbase + np] = e[n];
                ++np;
            }
        if (moved) {
            QEB_CUDA(c, cudaMemcpyAsync(B.idx.p, idx.data(), (size_t)notcnv * 4, cudaMemcpyHostToDevice, c->stream));
            const dim3 g((nvecx + 255) / 256, notcnv);
            if (cp)
                k_dav_gather_cols<true><<<g, tb, 0, c->stream>>>((double*)B.vc.p, (double*)B.wa.p, nvecx, nvecx, notcnv, (int*)B.idx.p);
            else
                k_dav_gather_cols<false><<<g, tb, 0, c->stream>>>((double*)B.vc.p, (double*)B.wa.p, nvecx, nvecx, notcnv, (int*)B.idx.p);
            QEB_CUDA(c, cudaGetLastError());
            QEB_CUDA(c, cudaMemcpyAsync(B.vc.p, B.wa.p, (size_t)nvecx * notcnv * el, cudaMemcpyDeviceToDevice, c->stream));
        }
        QEB_CUDA(c, cudaMemcpyAsync(ew_d + nbase, ew.data() + nbase, (size_t)notcnv * 8, cudaMemcpyHostToDevice, c->stream));
        const int nb1 = nbase;
        cplx* pnew = colp(psi, nb1);
        // ( H - e S ) |psi> of the unconverged roots (:340-400)
        if ((rc = qeb_dav_gemm_nn(c, kdim, sb, kdmx, nbase, B.vc.p, nvecx, notcnv, pnew, kdmx, 0))) return rc;
        k_dav_scale_cols<0><<<dim3((kdim + 255) / 256, notcnv), tb, 0, c->stream>>>(pnew, kdmx, kdim, notcnv, ew_d + nb1);
        QEB_CUDA(c, cudaGetLastError());
        if ((rc = qeb_dav_gemm_nn(c, kdim, hpsi, kdmx, nbase, B.vc.p, nvecx, notcnv, pnew, kdmx, 1))) return rc;
        if (npwx > npw) QEB_CUDA(c, launch_zero_pad_rows(pnew, npwx, npw, npol * notcnv, c->stream));  // :400-401
        // approximate inverse iteration, then normalisation (:407-460)
        if ((rc = qeb_g_psi(c, npwx, npw, notcnv, npol, pnew, ew_d + nb1))) return rc;
        if ((rc = qeb_col_dot(c, kdim, notcnv, pnew, kdmx, pnew, kdmx, ew_d))) return rc;  // npol = 2: pad rows are zero (:428-434)
        k_dav_scale_cols<1><<<dim3((kdim + 255) / 256, notcnv), tb, 0, c->stream>>>(pnew, kdmx, kdim, notcnv, ew_d);
        QEB_CUDA(c, cudaGetLastError());
        if (!cp && c->gstart == 2) QEB_CUDA(c, launch_zero_im_g0(pnew, npwx, notcnv, c->stream));  // regterg.f90:375
        // H and S on the new vectors (:465-467)
        if ((rc = apply_h(pnew, colp(hpsi, nb1), notcnv))) return rc;
        nhpsi += notcnv;
        if (uspp && (rc = apply_s(pnew, colp(spsi, nb1), notcnv))) return rc;
        // rows nb1.. of the reduced matrices (:478-499), then the Hermitian fill (:520-533)
        const int tot = nbase + notcnv;
        if ((rc = qeb_dav_gemm_cn(c, kdim, colp(hpsi, nb1), kdmx, notcnv, psi, kdmx, tot, red_at(B.hc, nb1, 0), nvecx))) return rc;
        if ((rc = qeb_dav_gemm_cn(c, kdim, colp(sb, nb1), kdmx, notcnv, psi, kdmx, tot, red_at(B.sc, nb1, 0), nvecx))) return rc;
        nbase = tot;
        QEB_CUDA(c, symmetrise(nb1, nbase));
        if ((rc = dav_diaghg(c, B, nbase, nvec, nvecx, cp, ew.data()))) return rc;  // :545
        // convergence (:556-575)
        notcnv = 0;
        for (int i = 0; i < nvec; ++i) {
            conv[i] = std::fabs(ew[i] - e[i]) < (btype[i] == 1 ? ethr : empty_ethr);
            if (!conv[i]) ++notcnv;
            e[i] = ew[i];
        }
        if (notcnv == 0 || nbase + notcnv > nvecx || dav_iter == 20) {  // :576-669
            if ((rc = qeb_dav_gemm_nn(c, kdim, psi, kdmx, nbase, B.vc.p, nvecx, nvec, evc_d, kdmx, 0))) return rc;  // :588
            if (notcnv == 0 || dav_iter == 20) break;
            // refresh psi, H psi, S psi (:600-632); the columns nvec.. of psi are the scratch, as in the reference
            QEB_CUDA(c, cudaMemcpyAsync(psi, evc_d, col * nvec, cudaMemcpyDeviceToDevice, c->stream));
            if (uspp) {
                if ((rc = qeb_dav_gemm_nn(c, kdim, spsi, kdmx, nbase, B.vc.p, nvecx, nvec, colp(psi, nvec), kdmx, 0))) return rc;
                QEB_CUDA(c, cudaMemcpy2DAsync(spsi, col, colp(psi, nvec), col, (size_t)kdim * 16, nvec, cudaMemcpyDeviceToDevice, c->stream));
            }
            if ((rc = qeb_dav_gemm_nn(c, kdim, hpsi, kdmx, nbase, B.vc.p, nvecx, nvec, colp(psi, nvec), kdmx, 0))) return rc;
            QEB_CUDA(c, cudaMemcpy2DAsync(hpsi, col, colp(psi, nvec), col, (size_t)kdim * 16, nvec, cudaMemcpyDeviceToDevice, c->stream));
            nbase = nvec;
            QEB_CUDA(c, cudaMemcpyAsync(ew_d, e, (size_t)nvec * 8, cudaMemcpyHostToDevice, c->stream));
            const dim3 g((nbase + 255) / 256, nbase);
            if (cp)
                k_dav_reset<true><<<g, tb, 0, c->stream>>>((double*)B.hc.p, (double*)B.sc.p, (double*)B.vc.p, nvecx, nbase, ew_d);
            else
                k_dav_reset<false><<<g, tb, 0, c->stream>>>((double*)B.hc.p, (double*)B.sc.p, (double*)B.vc.p, nvecx, nbase, ew_d);
            QEB_CUDA(c, cudaGetLastError());
            QEB_CUDA(c, cudaStreamSynchronize(c->stream));  // e (host) was the source of an async copy
        }
    }
    if (!evc_dev) QEB_CUDA(c, cudaMemcpyAsync(evc, evc_d, col * nvec, cudaMemcpyDeviceToHost, c->stream));
    QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    *notcnv_out = notcnv;
    *dav_iter_out = dav_iter;
    *nhpsi_out = nhpsi;
    return 0;
}

void qeb_solver_release(qeb_ctx* c)
{
    if (!c->solver) return;
    DavBufs& B = c->solver->B;
    for (DevBuf* b : {&B.psi, &B.hpsi, &B.spsi, &B.hc, &B.sc, &B.vc, &B.wa, &B.wb, &B.work, &B.ew, &B.evc_stage, &B.idx, &B.info})
        b->release();
    if (B.handle && solver_api().ok) solver_api().Destroy(B.handle);
    delete c->solver;
    c->solver = nullptr;
}

extern "C" {

int qeb_cegterg(qeb_ctx* c, int npw, int npwx, int nvec, int nvecx, int npol, void* evc, double ethr, int uspp, double* e,
                const int* btype, int* notcnv, int lrot, int* dav_iter, int* nhpsi)
{
    if (c && c->gamma) QEB_FAIL(c, 117, "qeb_cegterg: the grid is gamma_only (use qeb_regterg)");
    return dav_solve(c, npw, npwx, nvec, nvecx, npol, evc, ethr, uspp, e, btype, notcnv, lrot, dav_iter, nhpsi, "qeb_cegterg");
}

int qeb_regterg(qeb_ctx* c, int npw, int npwx, int nvec, int nvecx, void* evc, double ethr, int uspp, double* e, const int* btype,
                int* notcnv, int lrot, int* dav_iter, int* nhpsi)
{
    if (c && !c->gamma) QEB_FAIL(c, 118, "qeb_regterg: the grid is not gamma_only (use qeb_cegterg)");
    return dav_solve(c, npw, npwx, nvec, nvecx, 1, evc, ethr, uspp, e, btype, notcnv, lrot, dav_iter, nhpsi, "qeb_regterg");
}

}  // extern "C"
