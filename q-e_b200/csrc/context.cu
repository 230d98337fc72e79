// libqeb200, translation unit 1 of 6: the context behind the C ABI (include/qeb200.h) - lifetime, the setters that replace
// the Fortran module state h_psi reads, table upload, workspace sizing, and the device-side producers of the
// per-k-point state (g2_kin, init_us_2). The operators live in operators.cu, the host-array pipeline in pipeline.cu,
// CUDA-graph replay / multi-k in graphs.cu, NCCL / peer mapping / band groups in multigpu.cu, the Davidson pieces in
// dav_blocks.cu and solver.cu.
#include "context.h"

extern "C" {

// returns a device pointer for `p` (copying host data into `stage` if needed)
int stage_in(qeb_ctx* c, const void* p, size_t bytes, DevBuf& stage, const void** out)
{
    if (is_device_ptr(p)) {
        *out = p;
        return 0;
    }
    QEB_CUDA(c, stage.ensure(bytes));
    QEB_CUDA(c, cudaMemcpyAsync(stage.p, p, bytes, cudaMemcpyHostToDevice, c->stream));
    *out = stage.p;
    return 0;
}


const char* qeb_version(void) { return "qeb200 0.1 (sm_100a; FP64 DMMA projectors; radix-2/3/5 pencil FFT)"; }

int qeb_device_count(void)
{
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return ndev;
}

int qeb_create(qeb_ctx** out, int device, void* stream)
{
    if (!out) return 1;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) return 2;  // no CPU fallback
    if (device < 0 || device >= ndev) return 3;
    if (cudaSetDevice(device) != cudaSuccess) return 4;
    qeb_ctx* c = new qeb_ctx();
    c->device = device;
    if (stream) {
        c->stream = (cudaStream_t)stream;
    } else {
        if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) {
            delete c;
            return 5;
        }
        c->own_stream = true;
    }
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, device);
    c->num_sms = prop.multiProcessorCount;
    c->max_smem = prop.sharedMemPerBlockOptin;
    c->max_smem_sm = prop.sharedMemPerMultiprocessor;
    if (vloc_init_attrs(c->max_smem) != cudaSuccess) {
        delete c;
        return 6;
    }
    if (const char* e = getenv("QEB_BATCH")) c->batch_user = atoi(e);
    if (const char* e = getenv("QEB_VLOC_STREAMS")) c->vloc_nstreams = std::max(1, std::min(4, atoi(e)));
    for (int i = 1; i < c->vloc_nstreams; ++i)
        if (cudaStreamCreateWithFlags(&c->vstream[i], cudaStreamNonBlocking) != cudaSuccess) {
            delete c;
            return 7;
        }
    for (int i = 0; i < 5; ++i)
        if (cudaEventCreateWithFlags(&c->vev[i], cudaEventDisableTiming) != cudaSuccess) {
            delete c;
            return 7;
        }
    if (const char* e = getenv("QEB_PIPELINE_CHUNK")) c->pipeline_chunk = atoi(e);
    *out = c;
    return 0;
}

int qeb_destroy(qeb_ctx* c)
{
    if (!c) return 0;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    DevBuf* bufs[] = {&c->g2kin_d, &c->vkb_own, &c->tab_d, &c->plan_d, &c->g1_d, &c->g2_d, &c->v_d, &c->vp_d, &c->vpl_d,
                      &c->revx_d, &c->full_plan_d, &c->nlint_d, &c->deeq_d, &c->becp_d, &c->ps_d, &c->part_d,
                      &c->psi_stage, &c->hpsi_stage, &c->beta_stage, &c->misc_stage, &c->tabp_d, &c->planp_d,
                      &c->dist_tab_d, &c->sbuf_d, &c->rbuf_d, &c->qq_d, &c->us_int_d, &c->tab_beta_d, &c->us_scratch_d,
                      &c->us_in_d, &c->gv_g_d, &c->gv_mill_d, &c->igk_dev_d, &c->st_eig_d, &c->st_int_d, &c->gpsi_d, &c->dav_small_d, &c->rho_d, &c->wts_d, &c->vnc_d, &c->dnc_d,
                      &c->qnc_d, &c->nc_rho_d, &c->nc_comb_d, &c->nc_w_d, &c->nc_w2_d};
    for (void* m : c->p2p_opened) cudaIpcCloseMemHandle(m);
    c->p2p_opened.clear();
    qeb_solver_release(c);
    for (auto& g : c->graphs)
        if (g.exec) cudaGraphExecDestroy(g.exec);
    c->graphs.clear();
    for (auto& g : c->mgraphs)
        if (g.exec) cudaGraphExecDestroy(g.exec);
    c->mgraphs.clear();
    for (int j = 0; j < 2; ++j)
        if (c->mev[j]) cudaEventDestroy(c->mev[j]);
    for (DevBuf* b : bufs) b->release();
    c->p2p_flags_d.release();
    c->p2p_ptrs_d.release();
    if ((c->nccl_owned || c->nccl_rg_owned) && nccl_api().ok) {
        if (c->nccl && c->nccl_owned) nccl_api().CommDestroy(c->nccl);
        if (c->nccl_rg && c->nccl_rg_owned) nccl_api().CommDestroy(c->nccl_rg);
    }
    for (cudaEvent_t e : c->ev) cudaEventDestroy(e);
    for (cudaEvent_t e : c->ev_sub) cudaEventDestroy(e);
    for (cudaEvent_t e : c->ev_tail) cudaEventDestroy(e);
    for (int i = 0; i < 5; ++i)
        if (c->vev[i]) cudaEventDestroy(c->vev[i]);
    for (int i = 1; i < 4; ++i)
        if (c->vstream[i]) cudaStreamDestroy(c->vstream[i]);
    for (auto& r : c->hostreg) cudaHostUnregister((void*)r.p);
    if (c->s_in) cudaStreamDestroy(c->s_in);
    if (c->s_out) cudaStreamDestroy(c->s_out);
    if (c->own_stream) cudaStreamDestroy(c->stream);
    delete c;
    return 0;
}

const char* qeb_last_error(const qeb_ctx* c) { return c ? c->err.c_str() : "null context"; }

int qeb_set_batch(qeb_ctx* c, int nboxes)
{
    if (!c || nboxes < 0) return 1;
    ++c->state_epoch;
    c->batch_user = nboxes;
    c->have_k = false;  // workspaces are re-sized by the next qeb_set_kpoint
    return 0;
}

int qeb_set_pipeline(qeb_ctx* c, int chunk_cols, int min_cols)
{
    if (!c || chunk_cols < 0 || min_cols < 1) return 1;
    c->pipeline_chunk = chunk_cols;
    c->pipeline_min_cols = min_cols;
    return 0;
}

int qeb_set_async(qeb_ctx* c, int async)
{
    if (!c) return 1;
    c->async = async != 0;
    return 0;
}

int qeb_pin_host(qeb_ctx* c, void* p, size_t bytes)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (!p || bytes == 0) QEB_FAIL(c, 110, "qeb_pin_host: bad arguments");
    for (auto& r : c->hostreg)
        if (r.p == (const char*)p) QEB_FAIL(c, 111, "qeb_pin_host: this address is already pinned (qeb_unpin_host first)");
    QEB_CUDA(c, cudaHostRegister(p, bytes, cudaHostRegisterDefault));
    c->hostreg.push_back({(const char*)p, bytes});
    return 0;
}

int qeb_unpin_host(qeb_ctx* c, void* p)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    for (size_t i = 0; i < c->hostreg.size(); ++i)
        if (c->hostreg[i].p == (const char*)p) {
            // copies queued on the pipeline streams may still read the range
            QEB_CUDA(c, cudaStreamSynchronize(c->stream));
            if (c->s_in) QEB_CUDA(c, cudaStreamSynchronize(c->s_in));
            if (c->s_out) QEB_CUDA(c, cudaStreamSynchronize(c->s_out));
            c->hostreg.erase(c->hostreg.begin() + i);
            QEB_CUDA(c, cudaHostUnregister(p));
            return 0;
        }
    QEB_FAIL(c, 112, "qeb_unpin_host: this address was not pinned with qeb_pin_host");
}

// plans for the x direction permutation of V and for the full-box FFT
static int setup_full_plans(qeb_ctx* c, int nr1, int nr2, int nr3)
{
    HostPlan hx, hy, hz;
    if (!make_host_plan(nr1, hx) || !make_host_plan(nr2, hy) || !make_host_plan(nr3, hz))
        QEB_FAIL(c, 13, "qeb_set_grid: FFT dimensions must factor into 2, 3, 5 (fft_support.f90:61-120)");
    c->nr1 = nr1;
    c->nr2 = nr2;
    c->nr3 = nr3;
    const size_t ntw = (size_t)nr1 + nr2 + nr3;
    QEB_CUDA(c, c->full_plan_d.ensure(ntw * 16 + ntw * 4));
    cplx* tw = (cplx*)c->full_plan_d.p;
    int* rv = (int*)(tw + ntw);
    const HostPlan* hp[3] = {&hx, &hy, &hz};
    FftPlan* fp[3] = {&c->fpx, &c->fpy, &c->fpz};
    size_t off = 0;
    for (int d = 0; d < 3; ++d) {
        QEB_CUDA(c, cudaMemcpyAsync(tw + off, hp[d]->tw.data(), hp[d]->tw.size() * 16, cudaMemcpyHostToDevice, c->stream));
        QEB_CUDA(c, cudaMemcpyAsync(rv + off, hp[d]->rev.data(), hp[d]->rev.size() * 4, cudaMemcpyHostToDevice, c->stream));
        fill_plan(*hp[d], tw + off, rv + off, *fp[d]);
        off += hp[d]->n;
    }
    QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    c->have_full = true;
    return 0;
}

int qeb_set_grid(qeb_ctx* c, int nr1, int nr2, int nr3, int nr1x, int nr2x, int nr3x, int ngm, const int* nl,
                 const int* nlm, int gamma_only)
{
    if (!c) return 1;
    ++c->state_epoch;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (nr1 < 1 || nr2 < 1 || nr3 < 1 || ngm < 1 || !nl) QEB_FAIL(c, 10, "qeb_set_grid: bad arguments");
    if (nr1x != nr1 || nr2x != nr2 || nr3x != nr3)
        QEB_FAIL(c, 11, "qeb_set_grid: padded leading dimensions (nr?x != nr?) are not supported");
    if (gamma_only && !nlm) QEB_FAIL(c, 12, "qeb_set_grid: gamma_only requires nlm");
    int rc = setup_full_plans(c, nr1, nr2, nr3);
    if (rc) return rc;
    c->ngm = ngm;
    c->gamma = gamma_only != 0;
    c->dist = false;
    c->nl.assign(nl, nl + ngm);
    if (c->gamma)
        c->nlm.assign(nlm, nlm + ngm);
    else
        c->nlm.clear();
    c->have_grid = true;
    c->have_k = false;
    c->have_v = false;
    return 0;
}

static int upload_tables(qeb_ctx* c, HostVlocTables& H, DevBuf& tab_d, DevBuf& plan_d, VlocTables& t, VlocLaunchCfg& cfg);
static int size_workspaces_dist(qeb_ctx* c);

// fft_type_descriptor of a slab ("R&G") distributed dffts: FFTXlib/src/fft_types.f90:35-185 (fields), :491-943
// (fft_type_set). nl/nlm are the rank's stick-major maps (fft_ggen.f90:77).
int qeb_set_grid_dist(qeb_ctx* c, int nr1, int nr2, int nr3, int nproc, int mype, const int* nr3p, const int* nsw,
                      const int* iss, const int* ismap, int nismap, int ngm, const int* nl, const int* nlm,
                      int gamma_only)
{
    if (!c) return 1;
    ++c->state_epoch;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (nr1 < 1 || nr2 < 1 || nr3 < 1 || nproc < 1 || mype < 0 || mype >= nproc || !nr3p || !nsw || !iss || !ismap ||
        ngm < 1 || !nl)
        QEB_FAIL(c, 14, "qeb_set_grid_dist: bad arguments");
    if (gamma_only && !nlm) QEB_FAIL(c, 12, "qeb_set_grid_dist: gamma_only requires nlm");
    int rc = setup_full_plans(c, nr1, nr2, nr3);
    if (rc) return rc;
    const int ncol = nr1 * nr2;
    DistHost& D = c->DH;
    D = DistHost();
    D.nproc = nproc;
    D.me = mype;
    int zsum = 0, ssum = 0;
    for (int p = 0; p < nproc; ++p) {
        if (nr3p[p] < 1) QEB_FAIL(c, 15, "qeb_set_grid_dist: a rank without planes (fft_types.f90:633-635)");
        if (nsw[p] < 0 || iss[p] < 0 || iss[p] + nsw[p] > nismap) QEB_FAIL(c, 15, "qeb_set_grid_dist: bad nsw/iss");
        D.nr3p.push_back(nr3p[p]);
        D.i0r3p.push_back(zsum);
        D.nsw.push_back(nsw[p]);
        D.prefix.push_back(ssum);
        zsum += nr3p[p];
        ssum += nsw[p];
    }
    if (zsum != nr3) QEB_FAIL(c, 15, "qeb_set_grid_dist: sum(nr3p) != nr3");
    D.nst_tot = ssum;
    // wave stick columns rank by rank; within a rank the exchange order is (x, y) ascending, which every rank can
    // derive from ismap alone
    c->col_owner.assign(ncol, -1);
    c->col_sidx.assign(ncol, -1);
    c->all_cols.clear();
    c->my_cols.clear();
    for (int p = 0; p < nproc; ++p) {
        std::vector<int> cols(ismap + iss[p], ismap + iss[p] + nsw[p]);
        for (int& v : cols) {
            v -= 1;
            if (v < 0 || v >= ncol || c->col_owner[v] >= 0) QEB_FAIL(c, 16, "qeb_set_grid_dist: bad ismap");
            c->col_owner[v] = p;
        }
        std::sort(cols.begin(), cols.end(), [nr1](int a, int b) {
            const int xa = a % nr1, xb = b % nr1;
            return xa != xb ? xa < xb : a < b;
        });
        for (int i = 0; i < (int)cols.size(); ++i) c->col_sidx[cols[i]] = i;
        c->all_cols.insert(c->all_cols.end(), cols.begin(), cols.end());
        if (p == mype) c->my_cols = cols;
    }
    // the rank's maps converted to box positions col + ncol*z, so that the table builder of the serial path applies
    const int my_off = iss[mype];
    auto conv = [&](int v) {
        const int s = (v - 1) / nr3, z = (v - 1) % nr3;
        if (v < 1 || my_off + s >= nismap) return 0;  // 0: invalid (only an error if igk refers to it)
        const int col = ismap[my_off + s] - 1;
        if (col < 0 || col >= ncol) return 0;
        return 1 + col + ncol * z;
    };
    c->nl.resize(ngm);
    for (int i = 0; i < ngm; ++i) c->nl[i] = conv(nl[i]);
    c->nlm.clear();
    if (gamma_only) {
        c->nlm.resize(ngm);
        for (int i = 0; i < ngm; ++i) c->nlm[i] = conv(nlm[i]);
    }
    c->ngm = ngm;
    c->gamma = gamma_only != 0;
    // plane tables: all wave sticks, this rank's z planes
    std::string e = build_vloc_tables(nr1, nr2, nr3p[mype], 0, nullptr, nullptr, nullptr, 0, c->Hp, 72 * 1024,
                                      c->all_cols.data(), (int)c->all_cols.size(), false, c->gamma);
    if (!e.empty()) QEB_FAIL(c, 17, "qeb_set_grid_dist: " + e);
    if (c->Hp.nst != D.nst_tot) QEB_FAIL(c, 17, "qeb_set_grid_dist: inconsistent stick count");
    if ((rc = upload_tables(c, c->Hp, c->tabp_d, c->planp_d, c->Tp, c->cfgp))) return rc;
    // exchange tables: owner and position-in-owner of every plane-ordered stick
    {
        const int nst = c->Hp.nst;
        std::vector<int> host((size_t)4 * nst + 3 * nproc + nr3);
        for (int sg = 0; sg < nst; ++sg) {
            const int col = c->Hp.stick_x[sg] + nr1 * c->Hp.stick_y[sg];
            host[sg] = c->col_owner[col];
            host[nst + sg] = c->col_sidx[col];
        }
        for (int p = 0; p < nproc; ++p) {
            host[2 * nst + p] = D.nsw[p];
            host[2 * nst + nproc + p] = D.prefix[p];
        }
        // fused exchange (VlocTables::peer_* / r_*): per plane-ordered stick the owner's prefix and stick count, per
        // plane its owner, per rank its first plane
        int* fx = host.data() + 2 * nst + 2 * nproc;
        for (int sg = 0; sg < nst; ++sg) {
            fx[sg] = D.prefix[host[sg]];
            fx[nst + sg] = D.nsw[host[sg]];
        }
        for (int p = 0; p < nproc; ++p) {
            fx[2 * nst + nr3 + p] = D.i0r3p[p];
            for (int z = D.i0r3p[p]; z < D.i0r3p[p] + D.nr3p[p]; ++z) fx[2 * nst + z] = p;
        }
        QEB_CUDA(c, c->dist_tab_d.ensure(host.size() * 4 + 16));
        QEB_CUDA(c, cudaMemcpyAsync(c->dist_tab_d.p, host.data(), host.size() * 4, cudaMemcpyHostToDevice, c->stream));
        QEB_CUDA(c, cudaStreamSynchronize(c->stream));
        const int* b = (const int*)c->dist_tab_d.p;
        c->DD.nproc = nproc;
        c->DD.me = mype;
        c->DD.nz_me = nr3p[mype];
        c->DD.nst_tot = nst;
        c->DD.nst_me = nsw[mype];
        c->DD.own = b;
        c->DD.sidx = b + nst;
        c->DD.nsw = b + 2 * nst;
        c->DD.prefix = b + 2 * nst + nproc;
        const int* fxd = b + 2 * nst + 2 * nproc;
        c->fx_rpref = fxd;
        c->fx_rnsw = fxd + nst;
        c->fx_zown = fxd + 2 * nst;
        c->fx_zi0 = fxd + 2 * nst + nr3;
    }
    c->dist = true;
    c->have_grid = true;
    c->have_k = false;
    c->have_v = false;
    return 0;
}

static int upload_tables(qeb_ctx* c, HostVlocTables& H, DevBuf& tab_d, DevBuf& plan_d, VlocTables& t, VlocLaunchCfg& cfg)
{
    // integer tables in one allocation
    std::vector<const std::vector<int>*> vs = {&H.xmap, &H.plane_s0, &H.plane_ns, &H.ymap, &H.grp_stick, &H.grp_nl,
                                               &H.grp_coef_off, &H.coef_ig, &H.coef_slot, &H.coef_slot2,
                                               &H.pl_ymapT, &H.plane_x};
    size_t tot = 0;
    std::vector<size_t> offs;
    for (auto v : vs) {
        offs.push_back(tot);
        tot += (v->size() + 3) & ~(size_t)3;
    }
    const size_t div_off = tot;
    tot += H.grp_div.size() * 3;
    QEB_CUDA(c, tab_d.ensure(tot * 4 + 16));
    std::vector<int> host(tot, 0);
    for (size_t i = 0; i < vs.size(); ++i)
        if (!vs[i]->empty()) memcpy(host.data() + offs[i], vs[i]->data(), vs[i]->size() * 4);
    if (!H.grp_div.empty()) memcpy(host.data() + div_off, H.grp_div.data(), H.grp_div.size() * 12);
    QEB_CUDA(c, cudaMemcpyAsync(tab_d.p, host.data(), tot * 4, cudaMemcpyHostToDevice, c->stream));
    const int* base = (const int*)tab_d.p;
    memset(&t, 0, sizeof(t));
    t.nr1 = H.nr1;
    t.nr2 = H.nr2;
    t.nr3 = H.nr3;
    t.nst = H.nst;
    t.nxw = H.nxw;
    t.npos = H.nr2 * H.nr3;
    t.xmap = base + offs[0];
    t.plane_s0 = base + offs[1];
    t.plane_ns = base + offs[2];
    t.ymap = base + offs[3];
    t.grp_stick = base + offs[4];
    t.grp_nl = base + offs[5];
    t.grp_coef_off = base + offs[6];
    t.coef_ig = base + offs[7];
    t.coef_slot = base + offs[8];
    t.coef_slot2 = H.gamma ? base + offs[9] : nullptr;
    t.grp_div = (const FastDiv*)(base + div_off);
    t.pl_ymapT = base + offs[10];
    t.pl_xact = base + offs[11];
    t.pl_ldp = H.pl_ldp;
    t.pl_v = nullptr;  // set by prepare_v when the plane kernel is usable
    t.TZ = H.TZ;
    t.ngrp = H.ngrp;
    t.TY = H.TY;
    t.nty = H.nty;
    t.divTY = make_fastdiv(H.TY);
    t.divTYlast = make_fastdiv(H.nr3 - (H.nty - 1) * H.TY);
    t.TX = H.TX;
    t.ntx = H.ntx;
    t.divTX = make_fastdiv(H.TX);
    t.divTXlast = make_fastdiv(t.npos - (H.ntx - 1) * H.TX);
    t.nsmax = H.nsmax;
    // plans
    const size_t ntw = (size_t)H.nr1 + H.nr2 + H.nr3;
    QEB_CUDA(c, plan_d.ensure(ntw * 20));
    cplx* tw = (cplx*)plan_d.p;
    int* rv = (int*)(tw + ntw);
    const HostPlan* hp[3] = {&H.px, &H.py, &H.pz};
    FftPlan* fp[3] = {&t.px, &t.py, &t.pz};
    size_t off = 0;
    for (int d = 0; d < 3; ++d) {
        if (hp[d]->tw.empty()) continue;  // plane tables carry no z plan
        QEB_CUDA(c, cudaMemcpyAsync(tw + off, hp[d]->tw.data(), hp[d]->tw.size() * 16, cudaMemcpyHostToDevice, c->stream));
        QEB_CUDA(c, cudaMemcpyAsync(rv + off, hp[d]->rev.data(), hp[d]->rev.size() * 4, cudaMemcpyHostToDevice, c->stream));
        fill_plan(*hp[d], tw + off, rv + off, *fp[d]);
        off += hp[d]->n;
    }
    QEB_CUDA(c, cudaStreamSynchronize(c->stream));  // host vectors go out of scope
    t.vp = (const double*)c->vp_d.p;
    cfg.smem_z = H.smem_z;
    cfg.smem_y = H.smem_y;
    cfg.smem_x = H.smem_x;
    cfg.threads_z = cfg.threads_y = cfg.threads_x = 256;
    cfg.smem_plane = H.smem_plane;
    cfg.max_smem = c->max_smem;
    cfg.max_smem_sm = c->max_smem_sm;
    cfg.num_sms = c->num_sms;
    if (H.smem_z > c->max_smem || H.smem_y > c->max_smem || H.smem_x > c->max_smem)
        QEB_FAIL(c, 21, "FFT dimension too large for the shared-memory pencil kernels");
    return 0;
}

// batch size and the G1/G2 workspaces of the single-GPU path
static int size_workspaces(qeb_ctx* c)
{
    HostVlocTables& H = c->H;
    VlocTables& t = c->T;
    t.g1_box = (size_t)H.nst * H.nr3;
    t.g1_zs = (size_t)H.nst;
    t.g2_box = (size_t)H.nxw * t.npos;
    int nb = c->batch_user;
    if (nb <= 0) {
        // enough boxes per launch to fill the SMs (measured on B200, tools/vloc_sweep.py: throughput saturates from
        // ~8 boxes of a large grid; keeping the planes L2-resident matters less than filling the machine)
        const size_t target = (size_t)192 << 20;
        nb = (int)(target / (t.g2_box * 16 + 1));
        if (nb < 8) nb = 8;
        if (nb > 64) nb = 64;
        // with the fused plane kernel G2 is not touched and the batch only has to fill the z-pass and plane grids:
        // measured at AUSURF112 (batch sweep 10..28): 20 boxes 26.55, 15 boxes 26.91 us/band
        if (H.smem_plane > 0 && H.smem_plane <= c->max_smem && vloc_plane_available(t) && nb < 20) nb = 20;
    }
    c->batch = nb;
    QEB_CUDA(c, c->g1_d.ensure(t.g1_box * 16 * nb * c->vloc_nstreams));
    QEB_CUDA(c, c->g2_d.ensure(t.g2_box * 16 * nb * c->vloc_nstreams));
    t.g1 = (cplx*)c->g1_d.p;
    t.g2 = (cplx*)c->g2_d.p;
    return 0;
}

// slab-distributed path: S (own sticks, all z), R (all sticks, own z, rank-blocked), G1 and G2 of the plane passes
static int size_workspaces_dist(qeb_ctx* c)
{
    const DistHost& D = c->DH;
    VlocTables& tz = c->T;
    VlocTables& tp = c->Tp;
    const size_t nst_me = (size_t)D.nsw[D.me], nz_me = (size_t)D.nr3p[D.me];
    if ((size_t)c->H.nst != nst_me)
        QEB_FAIL(c, 26, "qeb_set_kpoint: plane waves outside the rank's wave sticks (igk/nl do not match the descriptor)");
    tp.g1_box = (size_t)tp.nst * nz_me;
    tp.g1_zs = (size_t)tp.nst;
    tp.g2_box = (size_t)tp.nxw * tp.npos;
    int nb = c->batch_user;
    if (nb <= 0) {
        // enough boxes per launch to fill the SMs (measured: throughput saturates from ~8 boxes of a large grid,
        // L2 residency of the planes matters less), bounded by workspace memory
        // from the largest slab, so that every rank of the group picks the same batch (the exchange needs it)
        int nzmax = 1;
        for (int z : D.nr3p) nzmax = std::max(nzmax, z);
        const size_t target = (size_t)192 << 20;
        nb = (int)(target / ((size_t)tp.nxw * c->nr2 * nzmax * 16 + 1));
        if (nb < 8) nb = 8;
        if (nb > 64) nb = 64;
    }
    const size_t ns = (size_t)c->vloc_nstreams;
    if (c->p2p_exported && (nb != c->batch || nst_me * c->nr3 * 16 * nb * ns > c->sbuf_d.cap ||
                            tp.g1_box * 16 * nb * ns > c->rbuf_d.cap))
        QEB_FAIL(c, 27, "qeb_set_kpoint: the exchange buffers were exported to the peers (qeb_p2p_export) and cannot "
                        "change size; export after the k-point with the largest npw / before changing the batch");
    c->batch = nb;
    QEB_CUDA(c, c->sbuf_d.ensure(std::max<size_t>(16, nst_me * c->nr3 * 16 * nb * ns)));
    QEB_CUDA(c, c->rbuf_d.ensure(std::max<size_t>(16, tp.g1_box * 16 * nb * ns)));
    QEB_CUDA(c, c->g1_d.ensure(std::max<size_t>(16, tp.g1_box * 16 * nb * ns)));
    QEB_CUDA(c, c->g2_d.ensure(std::max<size_t>(16, tp.g2_box * 16 * nb * ns)));
    tz.g1 = (cplx*)c->sbuf_d.p;
    tz.g2 = nullptr;
    tz.g1_box = nst_me;  // [z][box][stick]: see VlocTables::g1_zs
    tp.g1 = (cplx*)c->g1_d.p;
    tp.g2 = (cplx*)c->g2_d.p;
    return 0;
}

int qeb_set_kpoint(qeb_ctx* c, int npw, int npwx, const int* igk, const double* g2kin, const void* vkb, int nkb,
                   int gstart)
{
    if (!c) return 1;
    ++c->state_epoch;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (!c->have_grid) QEB_FAIL(c, 20, "qeb_set_kpoint: qeb_set_grid has not been called");
    // g2kin == NULL: qeb_g2_kin fills it on the device; vkb == NULL with nkb > 0: qeb_init_us_2 builds it on the device
    if (npw < 1 || npwx < npw || nkb < 0) QEB_FAIL(c, 22, "qeb_set_kpoint: bad arguments");
    if (c->gamma && igk) {
        for (int i = 0; i < npw; ++i)
            if (igk[i] != i + 1) QEB_FAIL(c, 23, "qeb_set_kpoint: gamma_only expects igk(i) = i");
        igk = nullptr;
    }
    if (!c->gamma && !igk) QEB_FAIL(c, 24, "qeb_set_kpoint: igk required for k-points");
    std::string e = build_vloc_tables(c->nr1, c->nr2, c->nr3, npw, c->nl.data(), c->gamma ? c->nlm.data() : nullptr, igk,
                                      c->ngm, c->H, 72 * 1024, c->dist ? c->my_cols.data() : nullptr,
                                      c->dist ? (int)c->my_cols.size() : 0);
    if (!e.empty()) QEB_FAIL(c, 25, "qeb_set_kpoint: " + e);
    c->npw = npw;
    c->npwx = npwx;
    c->nkb = nkb;
    c->gstart = gstart;
    int rc = upload_tables(c, c->H, c->tab_d, c->plan_d, c->T, c->cfg);
    if (rc) return rc;
    if ((rc = c->dist ? size_workspaces_dist(c) : size_workspaces(c))) return rc;
    QEB_CUDA(c, c->g2kin_d.ensure((size_t)npw * 8));
    c->have_g2kin = g2kin != nullptr;
    if (g2kin)
        QEB_CUDA(c, cudaMemcpyAsync(c->g2kin_d.p, g2kin, (size_t)npw * 8,
                                    is_device_ptr(g2kin) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
    c->igk_h.clear();
    if (igk) c->igk_h.assign(igk, igk + npw);
    c->igk_dev_ok = false;
    if (nkb > 0 && !vkb) {
        c->vkb_d = nullptr;  // until qeb_init_us_2
    } else if (nkb > 0) {
        if (is_device_ptr(vkb)) {
            c->vkb_d = (const cplx*)vkb;  // borrowed: the caller keeps it alive (like the reference's module array)
        } else {
            QEB_CUDA(c, c->vkb_own.ensure((size_t)npwx * nkb * 16));
            QEB_CUDA(c, cudaMemcpyAsync(c->vkb_own.p, vkb, (size_t)npwx * nkb * 16, cudaMemcpyHostToDevice, c->stream));
            c->vkb_d = (const cplx*)c->vkb_own.p;
        }
    } else {
        c->vkb_d = nullptr;
    }
    QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    c->have_k = true;
    c->becp_m = 0;
    return 0;
}

int prepare_v(qeb_ctx* c, const double* v)
{
    // distributed: v holds the rank's own z planes (dffts%nnr points of the slab, x fastest)
    const int nz = c->dist ? c->DH.nr3p[c->DH.me] : c->nr3;
    const size_t nnr = (size_t)c->nr1 * c->nr2 * nz;
    const void* vd = nullptr;
    int rc = stage_in(c, v, nnr * 8, c->v_d, &vd);
    if (rc) return rc;
    QEB_CUDA(c, c->vp_d.ensure(nnr * 8));
    const double scale = 1.0 / ((double)c->nr1 * c->nr2 * c->nr3);
    QEB_CUDA(c, vloc_prepare_v((const double*)vd, (double*)c->vp_d.p, c->fpx.rev, c->nr1, c->nr2, nz, scale, c->stream));
    c->T.vp = (const double*)c->vp_d.p;
    c->Tp.vp = (const double*)c->vp_d.p;
    // fused plane pass (serial descriptor): the potential again, permuted for the shared-memory plane
    c->T.pl_v = nullptr;
    if (!c->dist && c->cfg.smem_plane > 0 && c->cfg.smem_plane <= c->max_smem && vloc_plane_available(c->T)) {
        QEB_CUDA(c, c->vpl_d.ensure(nnr * 8));
        QEB_CUDA(c, vloc_prepare_v_plane((const double*)vd, (double*)c->vpl_d.p, c->T.px.rev, c->T.py.rev, c->nr1, c->nr2,
                                         nz, scale, c->stream));
        c->T.pl_v = (const double*)c->vpl_d.p;
    }
    c->have_v = true;
    return 0;
}

int qeb_set_vloc(qeb_ctx* c, const double* vrs)
{
    if (!c) return 1;
    ++c->state_epoch;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (!c->have_grid) QEB_FAIL(c, 30, "qeb_set_vloc: qeb_set_grid has not been called");
    if (!vrs) QEB_FAIL(c, 31, "qeb_set_vloc: null potential");
    int rc = prepare_v(c, vrs);
    if (rc) return rc;
    QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    return 0;
}

// v(dffts%nnr, 4) of vloc_psi_nc (PW/src/vloc_psi_acc.f90:291): four consecutive arrays, each permuted like vp
int prepare_v_nc(qeb_ctx* c, const double* v4)
{
    const size_t nnr = (size_t)c->nr1 * c->nr2 * c->nr3;
    const void* vd = nullptr;
    int rc = stage_in(c, v4, 4 * nnr * 8, c->v_d, &vd);
    if (rc) return rc;
    QEB_CUDA(c, c->vnc_d.ensure(4 * nnr * 8));
    const double scale = 1.0 / ((double)c->nr1 * c->nr2 * c->nr3);
    for (int k = 0; k < 4; ++k)
        QEB_CUDA(c, vloc_prepare_v((const double*)vd + k * nnr, (double*)c->vnc_d.p + k * nnr, c->fpx.rev, c->nr1, c->nr2,
                                   c->nr3, scale, c->stream));
    c->T.vnc = (const double*)c->vnc_d.p;
    c->have_vnc = true;
    return 0;
}

int qeb_set_vloc_nc(qeb_ctx* c, const double* v4)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (!c->have_grid) QEB_FAIL(c, 30, "qeb_set_vloc_nc: qeb_set_grid has not been called");
    if (!v4) QEB_FAIL(c, 31, "qeb_set_vloc_nc: null potential");
    if (c->dist || c->gamma) QEB_FAIL(c, 32, "qeb_set_vloc_nc: serial k-point descriptor only (noncollinear has no gamma tricks)");
    int rc = prepare_v_nc(c, v4);
    if (rc) return rc;
    QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    return 0;
}

int qeb_set_nonlocal(qeb_ctx* c, int nat, int nhm, const int* nh_of_atom, const int* ofsbeta, const double* deeq)
{
    if (!c) return 1;
    ++c->state_epoch;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (nat < 0 || nhm < 0) QEB_FAIL(c, 40, "qeb_set_nonlocal: bad arguments");
    if (nat > 0 && (!nh_of_atom || !ofsbeta || !deeq)) QEB_FAIL(c, 40, "qeb_set_nonlocal: null arrays");
    int nkb_max = 0;
    for (int a = 0; a < nat; ++a) {
        if (nh_of_atom[a] < 0 || nh_of_atom[a] > nhm || ofsbeta[a] < 0) QEB_FAIL(c, 41, "qeb_set_nonlocal: bad nh/ofsbeta");
        nkb_max = std::max(nkb_max, ofsbeta[a] + nh_of_atom[a]);
    }
    std::vector<int> host((size_t)nkb_max + 2 * (size_t)nat + 4, -1);
    for (int a = 0; a < nat; ++a)
        for (int ih = 0; ih < nh_of_atom[a]; ++ih) host[ofsbeta[a] + ih] = a;
    for (int a = 0; a < nat; ++a) {
        host[nkb_max + a] = ofsbeta[a];
        host[nkb_max + nat + a] = nh_of_atom[a];
    }
    QEB_CUDA(c, c->nlint_d.ensure(host.size() * 4));
    QEB_CUDA(c, cudaMemcpyAsync(c->nlint_d.p, host.data(), host.size() * 4, cudaMemcpyHostToDevice, c->stream));
    c->row_atom_d = (const int*)c->nlint_d.p;
    c->ofs_d = c->row_atom_d + nkb_max;
    c->nh_d = c->ofs_d + nat;
    const size_t nd = (size_t)nhm * nhm * nat;
    QEB_CUDA(c, c->deeq_d.ensure(nd * 8 + 8));
    if (nd)
        QEB_CUDA(c, cudaMemcpyAsync(c->deeq_d.p, deeq, nd * 8,
                                    is_device_ptr(deeq) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
    QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    c->nat = nat;
    c->nhm = nhm;
    c->have_nl = true;
    c->nl_rows = nkb_max;
    return 0;
}

// ---- device-side producers of the per-k-point state ----------
int qeb_set_gvectors(qeb_ctx* c, int ngm, const double* g, const int* mill)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (ngm < 1 || !g) QEB_FAIL(c, 80, "qeb_set_gvectors: bad arguments");
    QEB_CUDA(c, c->gv_g_d.ensure((size_t)ngm * 24));
    QEB_CUDA(c, cudaMemcpyAsync(c->gv_g_d.p, g, (size_t)ngm * 24,
                                is_device_ptr(g) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
    if (mill) {
        QEB_CUDA(c, c->gv_mill_d.ensure((size_t)ngm * 12));
        QEB_CUDA(c, cudaMemcpyAsync(c->gv_mill_d.p, mill, (size_t)ngm * 12,
                                    is_device_ptr(mill) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
    }
    QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    c->gv_ngm = ngm;
    return 0;
}

static int ensure_igk_dev(qeb_ctx* c)
{
    if (c->igk_dev_ok || c->igk_h.empty()) return 0;
    QEB_CUDA(c, c->igk_dev_d.ensure(c->igk_h.size() * 4));
    QEB_CUDA(c, cudaMemcpyAsync(c->igk_dev_d.p, c->igk_h.data(), c->igk_h.size() * 4, cudaMemcpyHostToDevice, c->stream));
    QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    c->igk_dev_ok = true;
    return 0;
}

int qeb_g2_kin(qeb_ctx* c, const double* xk, double tpiba2, double qcutz, double ecfixed, double q2sigma, double* g2kin_out)
{
    if (!c) return 1;
    ++c->state_epoch;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (!c->have_k) QEB_FAIL(c, 81, "qeb_g2_kin: qeb_set_kpoint has not been called");
    if (!c->gv_ngm) QEB_FAIL(c, 82, "qeb_g2_kin: qeb_set_gvectors has not been called");
    if (!xk) QEB_FAIL(c, 83, "qeb_g2_kin: null xk");
    for (int v : c->igk_h)
        if (v < 1 || v > c->gv_ngm) QEB_FAIL(c, 84, "qeb_g2_kin: igk outside the G-vector list");
    if (c->igk_h.empty() && c->npw > c->gv_ngm) QEB_FAIL(c, 84, "qeb_g2_kin: npw exceeds the G-vector list");
    int rc = ensure_igk_dev(c);
    if (rc) return rc;
    const int* igk_d = c->igk_h.empty() ? nullptr : (const int*)c->igk_dev_d.p;
    QEB_CUDA(c, launch_g2_kin(c->npw, igk_d, (const double*)c->gv_g_d.p, xk[0], xk[1], xk[2], tpiba2, qcutz, ecfixed,
                              q2sigma, (double*)c->g2kin_d.p, c->stream));
    c->have_g2kin = true;
    if (g2kin_out)
        QEB_CUDA(c, cudaMemcpyAsync(g2kin_out, c->g2kin_d.p, (size_t)c->npw * 8,
                                    is_device_ptr(g2kin_out) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, c->stream));
    if (g2kin_out && !is_device_ptr(g2kin_out)) {
        QEB_CUDA(c, cudaStreamSynchronize(c->stream));
        return 0;
    }
    return finish(c);
}

int qeb_set_beta_tables(qeb_ctx* c, int nsp, int nhm, int nbetam, int lmaxkb, const int* nh, const int* nhtol,
                        const int* nhtolm, const int* indv, int nqx, double dq, const double* tab_beta)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (nsp < 1 || nhm < 0 || nbetam < 0 || nqx < 4 || !(dq > 0.0) || !nh) QEB_FAIL(c, 85, "qeb_set_beta_tables: bad arguments");
    if (lmaxkb > 6) QEB_FAIL(c, 85, "qeb_set_beta_tables: lmaxkb > 6 is not supported");
    if (nhm > 0 && (!nhtol || !nhtolm || !indv || !tab_beta)) QEB_FAIL(c, 85, "qeb_set_beta_tables: null tables");
    const int lmax2 = (lmaxkb + 1) * (lmaxkb + 1);
    for (int t = 0; t < nsp; ++t) {
        if (nh[t] < 0 || nh[t] > nhm) QEB_FAIL(c, 86, "qeb_set_beta_tables: nh out of range");
        for (int ih = 0; ih < nh[t]; ++ih) {
            const int k = t * nhm + ih;
            if (nhtolm[k] < 1 || nhtolm[k] > lmax2 || indv[k] < 1 || indv[k] > nbetam || nhtol[k] < 0)
                QEB_FAIL(c, 86, "qeb_set_beta_tables: nhtol/nhtolm/indv out of range");
        }
    }
    const size_t nt = (size_t)nhm * nsp;
    std::vector<int> host((size_t)nsp + 3 * nt + 4, 0);
    memcpy(host.data(), nh, (size_t)nsp * 4);
    if (nt) {
        memcpy(host.data() + nsp, nhtol, nt * 4);
        memcpy(host.data() + nsp + nt, nhtolm, nt * 4);
        memcpy(host.data() + nsp + 2 * nt, indv, nt * 4);
    }
    QEB_CUDA(c, c->us_int_d.ensure(host.size() * 4));
    QEB_CUDA(c, cudaMemcpyAsync(c->us_int_d.p, host.data(), host.size() * 4, cudaMemcpyHostToDevice, c->stream));
    const size_t ntab = (size_t)nqx * nbetam * nsp;
    QEB_CUDA(c, c->tab_beta_d.ensure(ntab * 8 + 8));
    if (ntab)
        QEB_CUDA(c, cudaMemcpyAsync(c->tab_beta_d.p, tab_beta, ntab * 8,
                                    is_device_ptr(tab_beta) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
    QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    const int* b = (const int*)c->us_int_d.p;
    c->US = UsppTables{nsp, nhm, nbetam, lmaxkb, nqx, dq, b, b + nsp, b + nsp + nt, b + nsp + 2 * nt,
                       (const double*)c->tab_beta_d.p};
    c->us_nh.assign(nh, nh + nsp);
    c->nsp = nsp;
    c->have_beta = true;
    return 0;
}

int qeb_set_structure(qeb_ctx* c, int nat, const int* ityp, const double* tau, int nr1, int nr2, int nr3,
                      const void* eigts1, const void* eigts2, const void* eigts3)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (nat < 1 || !ityp || !tau || nr1 < 1 || nr2 < 1 || nr3 < 1 || !eigts1 || !eigts2 || !eigts3)
        QEB_FAIL(c, 87, "qeb_set_structure: bad arguments");
    const size_t n1 = (size_t)(2 * nr1 + 1) * nat, n2 = (size_t)(2 * nr2 + 1) * nat, n3 = (size_t)(2 * nr3 + 1) * nat;
    QEB_CUDA(c, c->st_eig_d.ensure((n1 + n2 + n3) * 16));
    cplx* e = (cplx*)c->st_eig_d.p;
    const void* src[3] = {eigts1, eigts2, eigts3};
    const size_t cnt[3] = {n1, n2, n3};
    size_t off = 0;
    for (int d = 0; d < 3; ++d) {
        QEB_CUDA(c, cudaMemcpyAsync(e + off, src[d], cnt[d] * 16,
                                    is_device_ptr(src[d]) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
        off += cnt[d];
    }
    QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    c->st_nat = nat;
    c->st_nr1 = nr1;
    c->st_nr2 = nr2;
    c->st_nr3 = nr3;
    c->st_ityp.assign(ityp, ityp + nat);
    c->st_tau.assign(tau, tau + 3 * (size_t)nat);
    return 0;
}

int qeb_init_us_2(qeb_ctx* c, const double* q, double tpiba, void* vkb_out)
{
    if (!c) return 1;
    ++c->state_epoch;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (!c->have_k) QEB_FAIL(c, 88, "qeb_init_us_2: qeb_set_kpoint has not been called");
    if (!c->have_beta) QEB_FAIL(c, 88, "qeb_init_us_2: qeb_set_beta_tables has not been called");
    if (!c->st_nat) QEB_FAIL(c, 88, "qeb_init_us_2: qeb_set_structure has not been called");
    if (!c->gv_ngm || !c->gv_mill_d.p) QEB_FAIL(c, 88, "qeb_init_us_2: qeb_set_gvectors (with mill) has not been called");
    if (!q) QEB_FAIL(c, 88, "qeb_init_us_2: null q");
    if (c->US.lmaxkb < 0) return 0;  // init_us_2_acc.f90:73
    for (int v : c->igk_h)
        if (v < 1 || v > c->gv_ngm) QEB_FAIL(c, 84, "qeb_init_us_2: igk outside the G-vector list");
    // column offsets in the reference's order: all atoms of type 1, then type 2, ... (:126-159)
    const int nat = c->st_nat;
    std::vector<int> jkb0(nat, 0);
    int jkb = 0;
    for (int nt = 1; nt <= c->nsp; ++nt)
        for (int na = 0; na < nat; ++na) {
            if (c->st_ityp[na] < 1 || c->st_ityp[na] > c->nsp) QEB_FAIL(c, 89, "qeb_init_us_2: ityp out of range");
            if (c->st_ityp[na] == nt) {
                jkb0[na] = jkb;
                jkb += c->us_nh[nt - 1];
            }
        }
    if (jkb != c->nkb) QEB_FAIL(c, 89, "qeb_init_us_2: sum of nh over atoms differs from the nkb of qeb_set_kpoint");
    const double tpi = 2.0 * 3.14159265358979323846;
    std::vector<cplx> qph(nat);
    for (int na = 0; na < nat; ++na) {
        const double arg = (q[0] * c->st_tau[3 * na] + q[1] * c->st_tau[3 * na + 1] + q[2] * c->st_tau[3 * na + 2]) * tpi;
        qph[na] = mk(cos(arg), -sin(arg));
    }
    // small per-call arrays: ityp | jkb0 | qphase
    const size_t ib = ((size_t)2 * nat * 4 + 15) & ~(size_t)15;
    QEB_CUDA(c, c->st_int_d.ensure(ib + (size_t)nat * 16));
    std::vector<unsigned char> host(ib + (size_t)nat * 16, 0);
    memcpy(host.data(), c->st_ityp.data(), (size_t)nat * 4);
    memcpy(host.data() + (size_t)nat * 4, jkb0.data(), (size_t)nat * 4);
    memcpy(host.data() + ib, qph.data(), (size_t)nat * 16);
    QEB_CUDA(c, cudaMemcpyAsync(c->st_int_d.p, host.data(), host.size(), cudaMemcpyHostToDevice, c->stream));
    int rc = ensure_igk_dev(c);
    if (rc) return rc;
    const int* igk_d = c->igk_h.empty() ? nullptr : (const int*)c->igk_dev_d.p;
    const int npw = c->npw, npwx = c->npwx;
    const int lmax2 = (c->US.lmaxkb + 1) * (c->US.lmaxkb + 1);
    const size_t nscr = (size_t)npw * (lmax2 + (size_t)c->US.nsp * c->US.nbetam);
    QEB_CUDA(c, c->us_scratch_d.ensure(nscr * 8 + 8));
    double* ylm = (double*)c->us_scratch_d.p;
    double* vq = ylm + (size_t)npw * lmax2;
    QEB_CUDA(c, launch_us2_ylm_vq(npw, igk_d, (const double*)c->gv_g_d.p, q[0], q[1], q[2], tpiba, c->US, ylm, vq, c->stream));
    QEB_CUDA(c, c->vkb_own.ensure((size_t)npwx * c->nkb * 16));
    const cplx* e1 = (const cplx*)c->st_eig_d.p;
    const cplx* e2 = e1 + (size_t)(2 * c->st_nr1 + 1) * nat;
    const cplx* e3 = e2 + (size_t)(2 * c->st_nr2 + 1) * nat;
    const int* ityp_d = (const int*)c->st_int_d.p;
    QEB_CUDA(c, launch_us2_vkb(npw, npwx, igk_d, (const int*)c->gv_mill_d.p, nat, ityp_d, ityp_d + nat,
                               (const cplx*)((const char*)c->st_int_d.p + ib), c->st_nr1, c->st_nr2, c->st_nr3, e1, e2, e3,
                               c->US, ylm, vq, (cplx*)c->vkb_own.p, c->stream));
    c->vkb_d = (const cplx*)c->vkb_own.p;
    c->becp_m = 0;
    if (vkb_out)
        QEB_CUDA(c, cudaMemcpyAsync(vkb_out, c->vkb_own.p, (size_t)npwx * c->nkb * 16,
                                    is_device_ptr(vkb_out) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, c->stream));
    QEB_CUDA(c, cudaStreamSynchronize(c->stream));  // host staging vectors go out of scope
    return 0;
}

int qeb_vloc_kernels_per_batch(qeb_ctx* c)
{
    if (!c) return 0;
    if (c->dist) return 7;
    const char* e = getenv("QEB_PLANE");
    const bool off = (e && atoi(e) == 0) || (getenv("QEB_GENERIC") && atoi(getenv("QEB_GENERIC")));
    return (c->T.pl_v && !off) ? 3 : 5;
}

int qeb_get_stats(qeb_ctx* c, double* vloc_bytes_per_band, double* gemm_flops_per_band, int* nst, int* nxw, int* nboxes)
{
    if (!c || !c->have_k) return 1;
    const HostVlocTables& H = c->H;
    double fftb = 16.0 * (4.0 * H.nst * H.nr3 + 4.0 * (double)H.nxw * H.nr2 * H.nr3);
    if (c->dist)  // this rank's share: own sticks over all z, all active x planes over own z
        fftb = 16.0 * (4.0 * H.nst * H.nr3 + 4.0 * (double)c->Hp.nxw * c->Hp.nr2 * c->Hp.nr3);
    if (c->gamma) fftb *= 0.5;
    if (vloc_bytes_per_band) *vloc_bytes_per_band = fftb + 60.0 * c->npw;
    if (gemm_flops_per_band) *gemm_flops_per_band = (c->gamma ? 8.0 : 16.0) * (double)c->nkb * c->npw;
    if (nst) *nst = H.nst;
    if (nxw) *nxw = c->dist ? c->Hp.nxw : H.nxw;
    if (nboxes) *nboxes = c->batch;
    return 0;
}

}  // extern "C"
