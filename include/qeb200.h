/*
 * qeb200 - C ABI of the B200-native H|psi> path of PWscf.
 *
 * This is the drop-in boundary: a Fortran ISO_C_BINDING shim (fortran/qeb200_shim.f90,
 * INTEGRATION.md) keeps the reference's entry points and forwards to these
 * functions. Every pointer may be a HOST or a DEVICE pointer (detected with
 * cudaPointerGetAttributes); host data is staged through the context's own
 * device buffers. All arrays use the reference's layout: Fortran column-major,
 * COMPLEX(DP) = interleaved (re,im) doubles, index maps INT32 and 1-based.
 *
 * The reference keeps the state h_psi reads in Fortran modules; here it lives
 * in an explicit context whose setters mirror the cadence at which that module
 * state changes (file:line relative to the QE 7.5 tree):
 *   qeb_set_grid      after fft_type_init + fft_set_nl  (PW/src/data_structure.f90:74, Modules/recvec_subs.f90:309)
 *   qeb_set_kpoint    after g2_kin + init_us_2          (PW/src/c_bands.f90:105-109)
 *   qeb_set_vloc      after set_vrs                     (PW/src/electrons.f90:1038)
 *   qeb_set_nonlocal  after newd                        (PW/src/electrons.f90:1045)
 *
 * Return value: 0 on success, non-zero on error (the shim turns it into
 * errore(routine, qeb_last_error(ctx), code), UtilXlib/error_handler.f90:10).
 * There is no CPU fallback: without a CUDA device every compute call fails.
 */
#ifndef QEB200_H
#define QEB200_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct qeb_ctx qeb_ctx;

/* ---- lifetime ---------------------------------------------------------- */
/* One context per (process, device) - the analogue of one MPI rank's module
 * state (Modules/mp_world.f90:94 binds rank -> device). stream: a cudaStream_t
 * cast to void* (NULL = the context creates its own non-blocking stream and
 * synchronises it before every entry point returns, which reproduces the
 * reference's blocking OpenACC regions). */
int qeb_create(qeb_ctx** ctx, int device, void* stream);
/* number of CUDA devices visible to the process (0 without a driver): the shim binds rank -> device with
 * mod(node-local rank, qeb_device_count()), as Modules/mp_world.f90:87-94 does */
int qeb_device_count(void);
int qeb_destroy(qeb_ctx* ctx);
const char* qeb_last_error(const qeb_ctx* ctx);
/* version / build info string */
const char* qeb_version(void);
/* bands (k) or band pairs (Gamma) in flight per V_loc batch; the reference's
 * many_fft = 16 (Modules/control_flags.f90:285-288). 0 = automatic. */
int qeb_set_batch(qeb_ctx* ctx, int nboxes);
/* asynchronous mode: 1 = entry points return without synchronising the
 * stream (for callers that own the stream, e.g. the benchmark). */
int qeb_set_async(qeb_ctx* ctx, int async);

/* Host-buffer pipeline of qeb_h_psi: when psi and hpsi are HOST arrays with at least min_cols columns, the call is
 * split into chunks of chunk_cols columns (0 = automatic, about m/8 rounded to 64) whose H2D copy, kernels and D2H
 * copy overlap on three streams. Columns are independent (h_psi.f90:53-67), results are unchanged.
 * The copies only overlap the kernels when the caller's arrays are page-locked. The library never page-locks caller
 * memory on its own (it cannot see the caller free and re-allocate an array): the host brackets the lifetime of its
 * work arrays with qeb_pin_host / qeb_unpin_host (cudaHostRegister), e.g. right after cegterg's ALLOCATE(psi, hpsi)
 * and before the DEALLOCATE (KS_Solvers/Davidson/cegterg.f90:150-175, :680-700). Pageable arrays work, slower. */
int qeb_set_pipeline(qeb_ctx* ctx, int chunk_cols, int min_cols);
int qeb_pin_host(qeb_ctx* ctx, void* p, size_t bytes);
int qeb_unpin_host(qeb_ctx* ctx, void* p); /* synchronises the context's streams first; call before freeing p */

/* ---- state setters ----------------------------------------------------- */
/* FFT grid + G-vector maps of dffts (serial descriptor):
 *   fft_type_descriptor nr1,nr2,nr3,nr1x,nr2x,nr3x,nnr   FFTXlib/src/fft_types.f90:35-185
 *   nl, nlm (nlm NULL unless gamma_only)                 FFTXlib/src/fft_ggen.f90:28-108
 * nr?x must equal nr? (true on every non-ESSL build, fft_support.f90:40-54). */
int qeb_set_grid(qeb_ctx* ctx, int nr1, int nr2, int nr3, int nr1x, int nr2x, int nr3x, int ngm, const int* nl,
                 const int* nlm, int gamma_only);

/* Per-k-point state: npw = ngk(ik), npwx, igk_k(:,ik) (NULL for gamma_only:
 * igk(i)=i), g2kin(1:npw) (PW/src/g2_kin.f90:37-43), vkb(npwx,nkb)
 * (upflib/init_us_2_acc.f90:10), gstart (Modules/recvec_subs.f90:256-260).
 * vkb may be NULL when nkb == 0, or when qeb_init_us_2 builds it; g2kin may be NULL when qeb_g2_kin builds it. */
int qeb_set_kpoint(qeb_ctx* ctx, int npw, int npwx, const int* igk, const double* g2kin, const void* vkb, int nkb,
                   int gstart);

/* Local potential on the smooth grid for current_spin: vrs(1:nnr,current_spin)
 * (PW/src/set_vrs.f90), x-fastest box. */
int qeb_set_vloc(qeb_ctx* ctx, const double* vrs);

/* Nonlocal D matrices for current_spin: deeq(nhm,nhm,nat,current_spin) real
 * (upflib/uspp.f90:56-89), nh(ityp(na)) per atom and ofsbeta(na)
 * (0-based row offset into vkb/becp of atom na's projectors). */
int qeb_set_nonlocal(qeb_ctx* ctx, int nat, int nhm, const int* nh_of_atom, const int* ofsbeta, const double* deeq);

/* ---- the reference entry points ---------------------------------------- */
/* h_psi(lda,n,m,psi,hpsi)           PW/src/h_psi.f90:9-80 / h_psi_gpu.f90:9-84
 * hpsi = (g2kin + V_loc + V_NL) psi for m bands; rows n+1..lda of hpsi are
 * zeroed; Gamma: Im hpsi(G=0) = 0 when gstart == 2. Side effect kept from the
 * reference: the projections <beta|psi> of these m bands are
 * left in the context and can be read with qeb_get_becp (the shim copies them
 * into becmod::becp for the s_psi call that follows, PW/src/s_psi.f90:15). */
int qeb_h_psi(qeb_ctx* ctx, int lda, int n, int m, const void* psi, void* hpsi);

/* vloc_psi_k(lda,n,m,psi,v,hpsi)    PW/src/vloc_psi_acc.f90:148-266
 * hpsi (INOUT) += V_loc psi. v == NULL: use the potential of qeb_set_vloc. */
int qeb_vloc_psi_k(qeb_ctx* ctx, int lda, int n, int m, const void* psi, const double* v, void* hpsi);
/* vloc_psi_gamma(lda,n,m,psi,v,hpsi) PW/src/vloc_psi_acc.f90:9-145 */
int qeb_vloc_psi_gamma(qeb_ctx* ctx, int lda, int n, int m, const void* psi, const double* v, void* hpsi);
/* vloc_psi_nc(lda,n,m,psi,v,hpsi)    PW/src/vloc_psi_acc.f90:269-369, domag branch (:329-341)   
 * Noncollinear spinors with magnetisation: psi(lda*2,m), hpsi(lda,2,m) INOUT, v(dffts%nnr,4):
 *   hpsi_up += F[ psi_up(r) (v1+v4) + psi_dn(r) (v2 - i v3) ],  hpsi_dn += F[ psi_dn(r) (v1-v4) + psi_up(r) (v2 + i v3) ].
 * v == NULL: the potential of qeb_set_vloc_nc. Serial k-point descriptor only. The non-magnetic branch (:342-348,
 * v(:,1) on both components) is qeb_vloc_psi_k on 2m columns of leading dimension lda. */
int qeb_set_vloc_nc(qeb_ctx* ctx, const double* v4);
int qeb_vloc_psi_nc(qeb_ctx* ctx, int lda, int n, int m, const void* psi, const double* v4, void* hpsi);
/* h_psi for noncollinear spinors with magnetisation / spin-orbit: PW/src/h_psi.f90:138-148 (kinetic term on both
 * components), :184-190 (vloc_psi_nc), :225-229 -> calbec_nc (Modules/becmod.f90:575-616: one ZGEMM over m*npol
 * columns) and add_vuspsi_nc (PW/src/add_vuspsi.f90:174-234).   
 * qeb_set_nonlocal_nc: deeq_nc(nhm,nhm,nat,4) COMPLEX(DP) (host or device); nat, nhm, nh, ofsbeta are those of a
 * preceding qeb_set_nonlocal (whose real deeq is not used by this path).
 * qeb_h_psi_nc: psi(lda*2,m) -> hpsi(lda*2,m); rows n+1..lda of both components zeroed; leaves becp%nc(nkb,2,m) in
 * the context (qeb_get_becp with 2*m columns). Needs qeb_set_vloc_nc. Serial k-point descriptor only. */
int qeb_set_nonlocal_nc(qeb_ctx* ctx, const void* deeq_nc);
int qeb_h_psi_nc(qeb_ctx* ctx, int lda, int n, int m, const void* psi, void* hpsi);
/* s_psi_nc   PW/src/s_psi.f90:308-370: spsi(lda*2,m) = psi + vkb (qq becp%nc), becp%nc from the preceding qeb_h_psi_nc.
 * qeb_set_qq_nc: qq_nc(nhm,nhm,nat,4) COMPLEX(DP) per atom = qq_so(:,:,1:4,ityp(na)) with spin-orbit (:344-350), else
 * qq_at(:,:,na) in blocks 1 and 4 (:336-342). Without it qeb_s_psi_nc is the norm-conserving copy. */
int qeb_set_qq_nc(qeb_ctx* ctx, const void* qq_nc);
int qeb_s_psi_nc(qeb_ctx* ctx, int lda, int n, int m, const void* psi, void* spsi);

/* invfft / fwfft (kind, f, dfft, howmany)  FFTXlib/src/fft_interfaces.f90:19-80,
 * fft_fwinv.f90:10-216: in-place 3-D transform of howmany boxes f(nnr*howmany).
 * sign > 0: G->R, exp(+i), unscaled; sign < 0: R->G, exp(-i), scaled 1/(nr1 nr2 nr3).
 * kind: QEB_FFT_RHO full box, QEB_FFT_WAVE the reference's "reduced" transform
 * (only the wave sticks/planes carry data; the rest of the box is
 * unspecified after R->G, exactly like cfft3ds, fft_scalar.FFTW3.f90:498-694;
 * here the full transform is always computed, like cfft3d_gpu,
 * fft_scalar.cuFFT.f90:695-698). */
#define QEB_FFT_RHO 0
#define QEB_FFT_WAVE 1
int qeb_fft(qeb_ctx* ctx, int sign, int kind, void* f, int howmany);

/* calbec(npw,beta,psi,betapsi,nbnd)  Modules/becmod.f90:138-172 (k: :431-486,
 * Gamma: :274-334): becp(nkb,m) = beta(1:n,:)^H psi(1:n,1:m); Gamma: real
 * result 2*Re(...) - [gstart==2] beta(1,i) psi(1,j). beta (npwx,nkb) with the
 * npwx/gstart of the last qeb_set_kpoint; beta == NULL uses the context's vkb.
 * becp: complex(nkb,m) for k, real(nkb,m) for Gamma. */
int qeb_calbec(qeb_ctx* ctx, int n, const void* beta, int nkb, const void* psi, int lda, int m, void* becp);

/* qeb_calbec without any allocation. No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it). Same
 * arguments and result as qeb_calbec, but beta (NULL: the context's vkb), psi and becp must be DEVICE pointers (nothing
 * is staged), the context must not be slab-distributed, and the kernel's split-K partial sums live in a file-scope
 * device array of QEB_STATIC_SCRATCH_MAX_BYTES that every GPU loading the library holds until the process exits.
 * Limit: nkb*m*16 bytes (k; *8 Gamma) <= QEB_STATIC_SCRATCH_MAX_BYTES, i.e. nkb*m <= 2097152 complex projections;
 * beyond it the call fails with code 75 (the split count is lowered to fit below it). Calls no cudaMalloc / cudaFree or
 * allocating library routine. All calls on one device share the array: calls that overlap on the GPU race. */
#define QEB_STATIC_SCRATCH_MAX_BYTES (32u << 20)
int qeb_calbec_static(qeb_ctx* ctx, int n, const void* beta, int nkb, const void* psi, int lda, int m, void* becp);
size_t qeb_static_scratch_bytes(void); /* the size compiled into the library (== QEB_STATIC_SCRATCH_MAX_BYTES) */

/* add_vuspsi(lda,n,m,hpsi)  PW/src/add_vuspsi.f90:10-237:
 * hpsi(1:n,1:m) += vkb * (blockdiag(deeq) * becp). becp == NULL: use the
 * projections left in the context by the last qeb_h_psi / qeb_calbec. */
int qeb_add_vuspsi(qeb_ctx* ctx, int lda, int n, int m, void* hpsi, const void* becp);

/* becp of the last qeb_h_psi / qeb_calbec: (nkb,m) complex (k) or real (Gamma) */
int qeb_get_becp(qeb_ctx* ctx, void* becp, int m);

/* Augmentation charges of the S matrix: qq_at(nhm,nhm,nat) real (upflib/uspp.f90:56-89; zero for atoms whose
 * type is not ultrasoft/PAW), same nat/nhm/ofsbeta as the last qeb_set_nonlocal. Call after qeb_set_nonlocal
 * (the reference fills it once per run, upflib/init_us_1.f90). */
int qeb_set_qq(qeb_ctx* ctx, int nat, int nhm, const double* qq_at);

/* s_psi(lda,n,m,psi,spsi)  PW/src/s_psi.f90:10-262 (s_psi_gamma :183-238, s_psi_k :242-306):
 * spsi = psi (all lda rows), and with an ultrasoft S (qeb_set_qq called, nkb > 0)
 * spsi(1:n,1:m) += vkb * (blockdiag(qq_at) * becp), where becp are the projections of THESE m bands left in the
 * context by the qeb_h_psi / qeb_calbec that preceded (s_psi.f90:15: "Requires the products of psi with all
 * beta functions in array becp(nkb,m) (calculated in h_psi or by calbec)"). Without qeb_set_qq it is the copy
 * the reference does for norm-conserving runs (:120 "IF (nkb == 0 .OR. .NOT. okvan) RETURN"). */
int qeb_s_psi(qeb_ctx* ctx, int lda, int n, int m, const void* psi, void* spsi);

/* ---- device-side producers of the per-k-point state --------- */
/* With these the caller passes NULL for g2kin and/or vkb in qeb_set_kpoint and the arrays are built on the device,
 * where h_psi consumes them (a k-point change then uploads igk only, not the 16*npwx*nkb bytes of vkb).
 *
 * qeb_set_gvectors: gvect::g(3,ngm) in 2pi/alat units and gvect::mill(3,ngm) (Modules/recvec.f90), once per run
 * (mill may be NULL when only qeb_g2_kin is used). */
int qeb_set_gvectors(qeb_ctx* ctx, int ngm, const double* g, const int* mill);
/* g2_kin(ik)  PW/src/g2_kin.f90:10-62: g2kin(i) = |xk + g(:,igk(i))|^2 tpiba2, plus the modified kinetic
 * functional qcutz (1 + erf((g2kin - ecfixed)/q2sigma)) when qcutz > 0. Uses npw/igk of the last qeb_set_kpoint;
 * the result becomes the context's g2kin and is also copied to g2kin_out when that is not NULL. */
int qeb_g2_kin(qeb_ctx* ctx, const double* xk, double tpiba2, double qcutz, double ecfixed, double q2sigma,
               double* g2kin_out);
/* uspp / uspp_param / beta_mod module state read by init_us_2 (upflib/init_us_2_acc.f90:22-24, beta_mod.f90:14-28):
 * nh(nsp); nhtol, nhtolm, indv (nhm,nsp) with the reference's 1-based lm / beta indices; lmaxkb;
 * tab_beta(nqx,nbetam,nsp) and its spacing dq. Once per run (init_us_1 / init_tab_beta). */
int qeb_set_beta_tables(qeb_ctx* ctx, int nsp, int nhm, int nbetam, int lmaxkb, const int* nh, const int* nhtol,
                        const int* nhtolm, const int* indv, int nqx, double dq, const double* tab_beta);
/* ions_base ityp(nat) (1-based), tau(3,nat) in alat units, and the structure-factor phases
 * eigts1(-nr1:nr1,nat), eigts2, eigts3 of the DENSE grid (PW/src/struct_fact.f90:60-87); after every ionic step. */
int qeb_set_structure(qeb_ctx* ctx, int nat, const int* ityp, const double* tau, int nr1, int nr2, int nr3,
                      const void* eigts1, const void* eigts2, const void* eigts3);
/* init_us_2(npw,npwx,igk,q,nat,tau,ityp,tpiba,omega,nr1,nr2,nr3,eigts1,eigts2,eigts3,mill,g,vkb)
 * upflib/init_us_2_acc.f90:10-172 (ylmr2.f90, interp_beta): vkb(npwx,nkb) for k-vector q (2pi/alat) with the
 * npw/npwx/igk/nkb of the last qeb_set_kpoint; it becomes the context's vkb; vkb_out (host or device, may be NULL)
 * receives a copy. */
int qeb_init_us_2(qeb_ctx* ctx, const double* q, double tpiba, void* vkb_out);

/* sum_band, density part (PW/src/sum_band.f90:326-437 Gamma, :440-647 k; get_rho_k :650-671, get_rho_gamma
 * :674-696): rho(1:nnr) (INOUT, smooth grid, x fastest) += sum_b w(b) |psi_b(r)|^2 for the m bands of evc(lda,m),
 * w(b) = wg(b,ik)/omega; Gamma: two bands per transform, w1 Re^2 + w2 Im^2. Uses grid / k-point of the context.
 * (becsum, the augmentation part, and the k/spin loops stay in the caller.) */
int qeb_sum_band_rho(qeb_ctx* ctx, int lda, int n, int m, const void* evc, const double* w, double* rho);
/* The `noncolin` branch of the same loop (PW/src/sum_band.f90:555-575; get_rho_domag :699-729): evc(2*lda,m) spinors
 * (up in rows 1:lda, down in rows lda+1:2*lda, as evc(npwx*npol,nbnd)); rho(nnr,1) += w (|up(r)|^2 + |dw(r)|^2) and,
 * when domag != 0, rho(nnr,2:4) += w * (2 Re(conj(up) dw), 2 Im(conj(up) dw), |up|^2 - |dw|^2). rho is (nnr,4) when
 * domag != 0, (nnr,1) otherwise. Not for gamma_only grids. */
int qeb_sum_band_rho_nc(qeb_ctx* ctx, int lda, int n, int m, const void* evc, const double* w, double* rho, int domag);

/* ---- block-Davidson building blocks ------------------------ */
/* The work arrays of cegterg/regterg (psi, hpsi, spsi: npwx x nvecx) live on the device in a GPU run; these entry
 * points take DEVICE pointers for them and host or device pointers for the small reduced matrices. gamma_only and
 * gstart are those of the context (qeb_set_grid / qeb_set_kpoint).
 *
 * g_psi(lda,n,m,npol,psi,e)  PW/src/g_psi.f90:10-70 (TEST_NEW_PRECONDITIONING branch, the one the file defines):
 * psi(i,k) /= 0.5 (1 + x + sqrt(1 + (x-1)^2)), x = h_diag(i) - e(k) s_diag(i); h_diag/s_diag = g_psi_mod
 * (set once per k-point with qeb_set_g_psi_diag). npol = 2: psi(lda,npol,m), h_diag / s_diag (lda,npol), i.e.
 * qeb_set_g_psi_diag with n = lda*npol (:44-62). */
int qeb_set_g_psi_diag(qeb_ctx* ctx, int n, const double* h_diag, const double* s_diag);
int qeb_g_psi(qeb_ctx* ctx, int lda, int n, int m, int npol, void* psi, const double* e);
/* ew(k) = <a_k|b_k> over n plane waves: MYDDOT(2*npw, a_k, b_k) (cegterg.f90:421); Gamma: 2*dot - a(1)b(1) when
 * gstart == 2 (regterg.f90:358-361). */
int qeb_col_dot(qeb_ctx* ctx, int n, int m, const void* a, int lda, const void* b, int ldb, double* out);
/* hc/sc = psi^H hpsi (ZGEMM 'C','N', cegterg.f90:209-229, 478-499); Gamma: hr = 2 psi^T hpsi - psi(1) hpsi(1)
 * (DGEMM + DGER, regterg.f90:203-221). C(ka,mb) complex (k) / real (Gamma), leading dimension ldc. */
int qeb_dav_gemm_cn(qeb_ctx* ctx, int n, const void* A, int lda, int ka, const void* B, int ldb, int mb, void* C,
                    int ldc);
/* psi_new = psi vc (ZGEMM 'N','N', cegterg.f90:353-395, 588-632; Gamma: DGEMM with real B on the 2*npwx real view):
 * C(1:n,1:m) = A(1:n,1:k) B(1:k,1:m) (+ C when accumulate != 0). */
int qeb_dav_gemm_nn(qeb_ctx* ctx, int n, const void* A, int lda, int k, const void* B, int ldb, int m, void* C,
                    int ldc, int accumulate);

/* hs_psi(lda,n,m,psi,hpsi,spsi)  PW/src/hs_psi.f90:7-40;  hs_1psi(lda,n,psi,hpsi,spsi)  PW/src/hs_1psi.f90:7-60 (m = 1):
 * H and S applied to the same block (the callers are the CG / PPCG / ParO solvers); S re-uses the becp of H. */
int qeb_hs_psi(qeb_ctx* ctx, int lda, int n, int m, const void* psi, void* hpsi, void* spsi);

/* h_psi of nk k-points in one call: context k applies its Hamiltonian to psi[k] on its own stream and all streams are
 * joined before the call returns (the k loop of PW/src/c_bands.f90:95-150; BASELINE configs 1-2 are launch-bound, so
 * the k-points of a pool run concurrently). Repeated device-pointer calls of a context with unchanged state and
 * arguments are replayed from a CUDA graph (one launch per h_psi; QEB_GRAPH=0 disables, =2 forces it for every size). */
int qeb_h_psi_multi(int nk, qeb_ctx* const* ctxs, const int* lda, const int* n, const int* m, const void* const* psi,
                    void* const* hpsi);

/* ---- device-resident block-Davidson solver ----------------- */
/* cegterg(h_psi_ptr, s_psi_ptr, uspp, g_psi_ptr, npw, npwx, nvec, nvecx, npol, evc, ethr, e, btype, notcnv, lrot,
 *         dav_iter, nhpsi)                                 KS_Solvers/Davidson/cegterg.f90:18-700
 * regterg(... same list without npol ...)                  KS_Solvers/Davidson/regterg.f90:18-640 (gamma_only grids)
 * The three EXTERNAL operators are the context's own (qeb_h_psi, qeb_s_psi when uspp != 0 - it needs qeb_set_qq for
 * ultrasoft overlaps -, qeb_g_psi with the diagonals of qeb_set_g_psi_diag). psi, hpsi, spsi (npwx x nvecx) and the
 * reduced matrices stay in device memory for the whole solve; `diaghg` is cusolverDnZhegvd / Dsygvd (libcusolver is
 * opened at run time, QEB_CUSOLVER_LIB overrides the name). evc (npwx x nvec, INOUT) may be a host or a device pointer;
 * e (nvec, OUT), btype (nvec, IN: 1 = occupied, threshold ethr; 0 = empty, threshold max(5 ethr, 1e-5)), notcnv,
 * dav_iter, nhpsi (OUT) are host data. npol = 2 (qeb_cegterg only): spinors evc(npwx*npol, nvec) with zero pad rows, the
 * operators are qeb_h_psi_nc / qeb_s_psi_nc, the products run over kdim = npwx*npol rows (cegterg.f90:131-142) and the
 * preconditioner diagonals are h_diag(npwx,npol), s_diag(npwx,npol) (qeb_set_g_psi_diag with n = npwx*npol). One band
 * group, serial diagonalisation; other cases return an error and the Fortran shim keeps the reference routine for them. */
int qeb_cegterg(qeb_ctx* ctx, int npw, int npwx, int nvec, int nvecx, int npol, void* evc, double ethr, int uspp, double* e,
                const int* btype, int* notcnv, int lrot, int* dav_iter, int* nhpsi);
int qeb_regterg(qeb_ctx* ctx, int npw, int npwx, int nvec, int nvecx, void* evc, double ethr, int uspp, double* e,
                const int* btype, int* notcnv, int lrot, int* dav_iter, int* nhpsi);

/* ---- multi-GPU (one process per GPU) ----------------------------------- */
/* divide_all(comm, m, m_start, m_end, recv_counts, displs)  UtilXlib/divide.f90:55-106.
 * 0-based start/end(exclusive) of group `me` of `nproc`. */
int qeb_divide_all(int m, int nproc, int me, int* start, int* end);
/* h_psi over band groups (PW/src/h_psi.f90:53-67): this rank applies H to its
 * own column block [start,end) of psi/hpsi (device or host pointers to the
 * FULL lda x m arrays); the all-gather of hpsi column blocks over
 * inter_bgrp_comm is done by the caller's NCCL communicator (the host-side
 * mirror q-e_b200/hpsi.py does it with torch.distributed; a Fortran caller
 * passes ncclComm through qeb_set_nccl). */
int qeb_h_psi_bgrp(qeb_ctx* ctx, int lda, int n, int m, const void* psi, void* hpsi, int nbgrp, int my_bgrp);
/* optional: ncclComm_t (as void*) of inter_bgrp_comm; when set, qeb_h_psi_bgrp
 * all-gathers the column blocks of a DEVICE hpsi itself. */
int qeb_set_nccl(qeb_ctx* ctx, void* nccl_comm, int nranks, int rank);

/* NCCL communicators by role. libnccl is bound at run time (dlopen: $QEB_NCCL_LIB, else an already loaded
 * libnccl.so.2, else the system one), so single-GPU use needs no NCCL.
 *   QEB_COMM_BGRP  inter_bgrp_comm: band groups (Modules/mp_bands.f90:56-123)
 *   QEB_COMM_RG    intra_bgrp_comm = dffts%comm: the R&G group of the distributed FFT (fft_types.f90:243-254)
 * qeb_set_nccl_comm borrows a communicator the caller created with the same libnccl; qeb_nccl_unique_id +
 * qeb_nccl_init create one inside the library (rank 0 makes the 128-byte id, the caller broadcasts it - MPI_Bcast in
 * a Fortran host, torch.distributed in the Python mirror - then every rank calls qeb_nccl_init). */
#define QEB_COMM_BGRP 0
#define QEB_COMM_RG 1
int qeb_set_nccl_comm(qeb_ctx* ctx, int role, void* nccl_comm, int nranks, int rank);
int qeb_nccl_unique_id(void* id128);
int qeb_nccl_init(qeb_ctx* ctx, int role, const void* id128, int nranks, int rank);

/* Slab ("R&G") distributed smooth grid: the per-rank fft_type_descriptor of a parallel run
 * (FFTXlib/src/fft_types.f90:35-185, filled by fft_type_set :491-943; slab decomposition = QE's default,
 * Modules/command_line_options.f90:29):
 *   nproc, mype           dfft%nproc, dfft%mype
 *   nr3p(nproc)           z planes per rank        nsw(nproc)  wave sticks per rank
 *   iss(nproc)            offset of each rank's sticks in ismap (0-based, as dfft%iss)
 *   ismap(nismap)         1-based plane position x + nr1x*(y-1)... of every stick, wave sticks first per rank
 *   nl, nlm (ngm)         the rank's stick-major maps nl = 1 + n3 + (isind-1)*nr3x (fft_ggen.f90:77)
 * After this call qeb_set_kpoint takes the rank's own npw/igk/g2kin/vkb rows, qeb_set_vloc the rank's z planes
 * of vrs (nr1*nr2*nr3p(mype) values), and qeb_h_psi / qeb_vloc_psi_* / qeb_calbec / qeb_add_vuspsi work on the
 * rank's rows: V_loc exchanges sticks <-> planes over the QEB_COMM_RG communicator (fft_scatter_2d.f90:38-266)
 * and calbec all-reduces becp over it (becmod.f90:395-399). */
int qeb_set_grid_dist(qeb_ctx* ctx, int nr1, int nr2, int nr3, int nproc, int mype, const int* nr3p, const int* nsw,
                      const int* iss, const int* ismap, int nismap, int ngm, const int* nl, const int* nlm,
                      int gamma_only);

/* Peer-to-peer form of the R&G exchange (one process per GPU of one NVLink/NVSwitch box): after qeb_set_grid_dist +
 * qeb_set_kpoint every rank calls qeb_p2p_export (publishes cudaIpc handles of its exchange buffers and flag words in
 * a QEB_P2P_BLOB_BYTES blob), the caller all-gathers the blobs in rank order (MPI_Allgather in a Fortran host,
 * torch.distributed in the Python mirror), every rank calls qeb_p2p_import, then a barrier. From then on
 * fft_scatter (fft_scatter_2d.f90:123,226) runs as copy-engine writes into the peers' buffers plus flag words, without
 * NCCL kernels; the becp all-reduce still uses the QEB_COMM_RG communicator. QEB_P2P=0 keeps the NCCL exchange. */
#define QEB_P2P_BLOB_BYTES 256
int qeb_p2p_export(qeb_ctx* ctx, void* blob);
int qeb_p2p_import(qeb_ctx* ctx, const void* blobs, int nranks);

/* ---- host-side descriptor builder (no GPU needed) ------------------------ */
/* For hosts that do not run the reference's own setup (benchmark, Python mirror, tests): the data of
 * fft_type_init + ggen + fft_set_nl + gk_sort + g2_kin for one k-point, and the slab distribution of the grid.
 *   grid dims     FFTXlib/src/fft_types.f90:1086-1197, fft_support.f90:61-165 (good_fft_order: factors 2,3,5)
 *   G-vectors     Modules/recvec_subs.f90:111-258 (|G|^2 order with eps8 ties, Gamma half space, gstart)
 *   nl / nlm      FFTXlib/src/fft_ggen.f90:28-108      igk / g2kin  PW/src/gk_sort.f90:9-98, g2_kin.f90:37-43
 *   distribution  FFTXlib/src/stick_base.f90:214-601 + fft_types.f90:491-943 (slab; pencil is not built)
 * at_cols: 3x3 row-major whose COLUMNS are a_1..a_3 in units of alat; xk in 2pi/alat (NULL = Gamma point);
 * ecutrho <= 0 means 4*ecutwfc. Cutoffs in Ry, alat in bohr. */
typedef struct qeb_desc qeb_desc;
int qeb_good_fft_order(int nr);
int qeb_desc_create(qeb_desc** d, const double* at_cols, double alat, double ecutwfc, double ecutrho,
                    const double* xk, int gamma_only);
int qeb_desc_destroy(qeb_desc* d);
const char* qeb_desc_error(const qeb_desc* d);
int qeb_desc_dims(const qeb_desc* d, int* nr1, int* nr2, int* nr3, int* ngm, int* npw, int* nst, int* gstart);
/* slab distribution over nproc ranks; afterwards the per-rank arrays are available */
int qeb_desc_distribute(qeb_desc* d, int nproc);
int qeb_desc_rank_dims(const qeb_desc* d, int rank, int* ngm, int* npw, int* gstart, int* nismap);
/* copies an array into `out` (caller-sized). rank < 0: the serial / global arrays; rank >= 0: that rank's. */
#define QEB_DESC_NL 0      /* int[ngm]    1-based: serial x-fastest box index; rank: stick-major */
#define QEB_DESC_NLM 1     /* int[ngm]    Gamma only */
#define QEB_DESC_IGK 2     /* int[npw]    1-based into the (rank's) G list */
#define QEB_DESC_G2KIN 3   /* double[npw] |k+G|^2 tpiba2 (Ry) */
#define QEB_DESC_MILL 4    /* int[3*ngm]  Miller indices (global only) */
#define QEB_DESC_GG 5      /* double[ngm] |G|^2 (global only) */
#define QEB_DESC_ISIND 6   /* int[nr1*nr2] wave-stick mask of the serial descriptor */
#define QEB_DESC_IPLW 7    /* int[nr1]    active x planes */
#define QEB_DESC_NR3P 8    /* int[nproc] */
#define QEB_DESC_NSW 9     /* int[nproc] */
#define QEB_DESC_NSP 10    /* int[nproc] */
#define QEB_DESC_ISS 11    /* int[nproc] */
#define QEB_DESC_ISMAP 12  /* int[nismap] */
#define QEB_DESC_IG_L2G 13 /* int[ngm_rank] 0-based index of the rank's G-vectors in the global list */
#define QEB_DESC_PW_L2G 14 /* int[npw_rank] 0-based row of the serial psi holding the rank's plane waves */
int qeb_desc_get(const qeb_desc* d, int what, int rank, void* out);

/* ---- introspection for benchmarks / tests ------------------------------ */
/* algorithmic HBM bytes per band of the V_loc path:
 * 16*(4*nst*nr3 + 4*nxw*nr2*nr3)[/2 Gamma] + 60*npw, and the table sizes */
int qeb_get_stats(qeb_ctx* ctx, double* vloc_bytes_per_band, double* gemm_flops_per_band, int* nst, int* nxw,
                  int* nboxes_batch);
/* kernels launched per V_loc batch with the current grid / potential: 3 when the fused plane kernel is active
 * (scatter+z, plane, z+gather), 5 otherwise; 7 on the slab-distributed path (pack / unpack around the exchange). */
int qeb_vloc_kernels_per_batch(qeb_ctx* ctx);
/* FP64 DMMA GEMM self-test entry used by the benchmark to time the projector
 * kernels alone: runs calbec + add_vuspsi on the context's vkb for m bands of
 * a device psi/hpsi. */
int qeb_nonlocal_only(qeb_ctx* ctx, int lda, int n, int m, const void* psi, void* hpsi);

#ifdef __cplusplus
}
#endif
#endif /* QEB200_H */
