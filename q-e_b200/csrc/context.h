// Internal header of libqeb200: the context behind the C ABI (include/qeb200.h) and the helpers every translation unit
// of the library shares. The context is the explicit form of the Fortran module state h_psi reads.
#pragma once
#include <cuda_runtime.h>
#include <algorithm>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>
#include "../../include/qeb200.h"
#include "distfft.h"
#include "fullfft.h"
#include "davidson.h"
#include "gemm.h"
#include "uspp.h"
#include "vloc.h"
#include "vloc_setup.h"

using namespace qeb;

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes)
    {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e == cudaSuccess) cap = bytes;
        return e;
    }
    void release()
    {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

inline bool is_device_ptr(const void* p)
{
    if (!p) return false;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

struct qeb_solver_state;  // solver.cu: work arrays of the device-resident Davidson solver
void qeb_solver_release(struct qeb_ctx* c);

struct qeb_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    bool async = false;
    int num_sms = 148;
    size_t max_smem = 0, max_smem_sm = 0;
    std::string err;

    // grid
    bool have_grid = false;
    int nr1 = 0, nr2 = 0, nr3 = 0, ngm = 0;
    bool gamma = false;
    std::vector<int> nl, nlm;
    // k-point
    bool have_k = false;
    int npw = 0, npwx = 0, nkb = 0, gstart = 1;
    DevBuf g2kin_d, vkb_own;
    bool have_g2kin = false;
    std::vector<int> igk_h;  // igk of qeb_set_kpoint (k-points), for qeb_g2_kin / qeb_init_us_2
    // uspp / beta_mod tables of qeb_set_beta_tables
    bool have_beta = false;
    int nsp = 0, us_nhm = 0, nbetam = 0, lmaxkb = -1, nqx = 0;
    double dq = 0.0;
    std::vector<int> us_nh;
    DevBuf us_int_d, tab_beta_d, us_scratch_d, us_in_d;
    UsppTables US;
    // G-vector list (qeb_set_gvectors) and structure factors (qeb_set_structure)
    int gv_ngm = 0;
    DevBuf gv_g_d, gv_mill_d, igk_dev_d;
    bool igk_dev_ok = false;
    int st_nat = 0, st_nr1 = 0, st_nr2 = 0, st_nr3 = 0;
    std::vector<int> st_ityp;
    std::vector<double> st_tau;
    DevBuf st_eig_d, st_int_d;
    // g_psi_mod::h_diag / s_diag (qeb_set_g_psi_diag) and small staging of the Davidson helpers
    DevBuf gpsi_d, dav_small_d, rho_d, wts_d;
    DevBuf nc_rho_d, nc_comb_d, nc_w_d, nc_w2_d;  // density_nc.cu: partial densities, combined columns, weights
    int gpsi_n = 0;
    const cplx* vkb_d = nullptr;
    // V_loc tables
    HostVlocTables H;
    DevBuf tab_d;  // all int tables
    DevBuf plan_d;  // twiddles + rev of the three plans
    VlocTables T;
    VlocLaunchCfg cfg;
    int batch_user = 0, batch = 1;
    DevBuf g1_d, g2_d;
    // batches are dealt round-robin to vloc_nstreams streams (each with its own G1/G2 slice), so that the
    // latency-bound z passes of one batch overlap the FP64-heavy x pass of another
    int vloc_nstreams = 2;
    cudaStream_t vstream[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t vev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    // potential
    bool have_v = false;
    DevBuf v_d, vp_d, vpl_d;
    DevBuf vnc_d;  // noncollinear 2x2 potential, four arrays in the vp layout
    bool have_vnc = false;
    DevBuf dnc_d;  // deeq_nc(nhm,nhm,nat,4) complex
    bool have_dnc = false;
    DevBuf qnc_d;  // per-atom 2x2-spin overlap (nhm,nhm,nat,4) complex: qq_so of the atom's type, or qq_at on the diagonal
    bool have_qnc = false;
    DevBuf revx_d;
    // full-box FFT plans
    DevBuf full_plan_d;
    FftPlan fpx, fpy, fpz;
    bool have_full = false;
    // nonlocal
    bool have_nl = false;
    int nat = 0, nhm = 0, nl_rows = 0;
    DevBuf nlint_d, deeq_d, qq_d;
    bool have_qq = false;
    const int *row_atom_d = nullptr, *ofs_d = nullptr, *nh_d = nullptr;
    DevBuf becp_d, ps_d, part_d;
    int becp_m = 0;
    // staging
    DevBuf psi_stage, hpsi_stage, beta_stage, misc_stage;
    // host-buffer pipeline (qeb_h_psi with host psi/hpsi): copy-in / copy-out streams, per-chunk events and the
    // caller's arrays page-locked once (the Davidson work arrays are re-used call after call)
    cudaStream_t s_in = nullptr, s_out = nullptr;
    std::vector<cudaEvent_t> ev;
    std::vector<cudaEvent_t> ev_sub;  // sub-chunks of the first pipeline chunk
    std::vector<cudaEvent_t> ev_tail;  // add_vuspsi sub-chunks of the last pipeline chunks
    struct HostReg { const char* p; size_t bytes; };
    std::vector<HostReg> hostreg;
    int pipeline_chunk = 0;  // 0 = automatic
    int pipeline_min_cols = 128;
    // nccl: inter_bgrp_comm (band groups) and intra_bgrp_comm (R&G distribution)
    void* nccl = nullptr;
    int nranks = 1, rank = 0;
    bool nccl_owned = false;
    void* nccl_rg = nullptr;
    bool nccl_rg_owned = false;
    // slab-distributed dffts (qeb_set_grid_dist)
    bool dist = false;
    DistHost DH;
    std::vector<int> my_cols, all_cols, col_owner, col_sidx;  // wave stick columns x + nr1*y
    HostVlocTables Hp;   // plane tables: every wave stick, the rank's own z planes
    VlocTables Tp;
    VlocLaunchCfg cfgp;
    DevBuf tabp_d, planp_d, dist_tab_d, sbuf_d, rbuf_d;
    DistDev DD;
    const int *fx_rpref = nullptr, *fx_rnsw = nullptr, *fx_zown = nullptr, *fx_zi0 = nullptr;  // fused exchange tables
    // peer-to-peer exchange (qeb_p2p_export / qeb_p2p_import)
    P2PState P2P;
    DevBuf p2p_flags_d, p2p_ptrs_d;
    bool p2p_exported = false;
    std::vector<void*> p2p_opened;  // cudaIpcOpenMemHandle mappings to close
    qeb_solver_state* solver = nullptr;
    // CUDA-graph replay of small h_psi calls (launch-bound k-points: SURVEY 7, BASELINE configs 1-2)
    struct GraphEntry {
        std::vector<unsigned long long> key;
        cudaGraphExec_t exec = nullptr;
        int seen = 0;
    };
    std::vector<GraphEntry> graphs;
    std::vector<GraphEntry> mgraphs;     // qeb_h_psi_multi: all k-points of a call in ONE graph (kept in the first context)
    cudaEvent_t mev[2] = {nullptr, nullptr};  // fork / join of this context's stream inside a multi-k capture
    unsigned long long state_epoch = 0;  // bumped by every setter that changes what the kernels are launched with
    int graph_mode = -1;                 // -1: automatic (small calls only), 0: off, 1: every device-pointer call
};

#define QEB_FAIL(ctx, code, msg)   \
    do {                           \
        (ctx)->err = (msg);        \
        return (code);             \
    } while (0)
#define QEB_CUDA(ctx, call)                                                                 \
    do {                                                                                    \
        cudaError_t e__ = (call);                                                           \
        if (e__ != cudaSuccess) {                                                           \
            (ctx)->err = std::string(#call) + ": " + cudaGetErrorString(e__);               \
            return 100 + (int)e__;                                                          \
        }                                                                                   \
    } while (0)

static inline bool env_flag(const char* name, int dflt)
{
    const char* e = getenv(name);
    return (e ? atoi(e) : dflt) != 0;
}

static inline int finish(qeb_ctx* c)
{
    if (!c->async) QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    return 0;
}

// ---- functions shared by the translation units of the library (context.cu: state and setters; operators.cu: the
// reference entry points; pipeline.cu: host-array pipeline; graphs.cu: CUDA-graph replay and multi-k; multigpu.cu: NCCL,
// peer mapping, band groups; dav_blocks.cu + solver.cu: Davidson) ----
extern "C" {
int stage_in(qeb_ctx* c, const void* p, size_t bytes, DevBuf& stage, const void** out);
int check_call(qeb_ctx* c, int lda, int n, int m, const void* psi, const void* hpsi, const char* who);
int prepare_v(qeb_ctx* c, const double* vrs);
int prepare_v_nc(qeb_ctx* c, const double* v4);
int run_vloc(qeb_ctx* c, int lda, int n, int m, const cplx* psi, cplx* hpsi, bool kinetic);
int ensure_becp(qeb_ctx* c, int nkb, int m);
int run_calbec(qeb_ctx* c, const cplx* beta, int npwx, int nkb, const cplx* psi, int lda, int n, int m, void* becp_out);
int run_addvus(qeb_ctx* c, int lda, int n, int m, cplx* hpsi, int col0 = 0, const double* dmat = nullptr);
int h_psi_cols(qeb_ctx* c, int lda, int n, int m, const cplx* psi_d, cplx* hpsi_d, bool do_vloc, bool do_nl, bool kinetic,
               int col0 = 0, int m_total = -1);
int h_psi_host_pipeline(qeb_ctx* c, int lda, int n, int m, const void* psi, void* hpsi);
bool graph_wanted(qeb_ctx* c, int n, int m);
int h_psi_graph(qeb_ctx* c, int lda, int n, int m, const cplx* psi_d, cplx* hpsi_d, bool* done);
}
