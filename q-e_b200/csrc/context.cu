// C ABI of the qeb200 library (include/qeb200.h): explicit context replacing
// the Fortran module state h_psi reads, host/device pointer
// handling, and the orchestration of the V_loc passes and projector GEMMs in
// the reference's order (PW/src/h_psi.f90:138-147 kinetic, :180/:220 V_loc,
// :229-236 nonlocal, :294-295 Gamma fix).
#include "context.h"

// returns a device pointer for `p` (copying host data into `stage` if needed)
static int stage_in(qeb_ctx* c, const void* p, size_t bytes, DevBuf& stage, const void** out)
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

extern "C" {

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
                      &c->qnc_d};
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

static int prepare_v(qeb_ctx* c, const double* v)
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
static int prepare_v_nc(qeb_ctx* c, const double* v4)
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

// sum_band, density part (PW/src/sum_band.f90:326-437 Gamma, :440-647 k): rho(r) += sum_b w_b |psi_b(r)|^2 for the
// m bands of evc, w_b = wg(b,ik)/omega. The reference loops wave_g2r + get_rho_k/get_rho_gamma per band (group);
// here: scatter + z pass, y pass, then one kernel that x-transforms every box of a batch and accumulates.
static int check_call(qeb_ctx* c, int lda, int n, int m, const void* psi, const void* hpsi, const char* who);
int qeb_sum_band_rho(qeb_ctx* c, int lda, int n, int m, const void* evc, const double* w, double* rho)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    int rc = check_call(c, lda, n, m, evc, rho, "qeb_sum_band_rho");
    if (rc) return rc;
    if (!w) QEB_FAIL(c, 97, "qeb_sum_band_rho: null weights");
    if (m == 0) return 0;
    // slab-distributed descriptor: rho holds the rank's own z planes (dffts%nr3p(mype) of them), as in the reference
    const size_t nnr = (size_t)c->nr1 * c->nr2 * (c->dist ? c->Tp.nr3 : c->nr3);
    const void* psi_d;
    if ((rc = stage_in(c, evc, (size_t)lda * m * 16, c->psi_stage, &psi_d))) return rc;
    const void* w_d;
    if ((rc = stage_in(c, w, (size_t)m * 8, c->wts_d, &w_d))) return rc;
    const bool rho_dev = is_device_ptr(rho);
    double* r = rho;
    if (!rho_dev) {
        QEB_CUDA(c, c->rho_d.ensure(nnr * 8));
        r = (double*)c->rho_d.p;
        QEB_CUDA(c, cudaMemcpyAsync(r, rho, nnr * 8, cudaMemcpyHostToDevice, c->stream));
    }
    VlocIO io;
    io.psi = (const cplx*)psi_d;
    io.hpsi = nullptr;
    io.g2kin = nullptr;
    io.lda = lda;
    io.n = n;
    io.nbands = m;
    io.gamma = c->gamma ? 1 : 0;
    const int nboxes = c->gamma ? (m + 1) / 2 : m;
    for (int b0 = 0; b0 < nboxes; b0 += c->batch) {  // one stream: batches add to the same rho
        const int nb = std::min(c->batch, nboxes - b0);
        io.band0 = c->gamma ? 2 * b0 : b0;
        if (!c->dist) {
            QEB_CUDA(c, vloc_launch_density(c->T, io, c->cfg, nb, (const double*)w_d, r, c->stream));
            continue;
        }
        // wave_g2r on the distributed grid (fft_parallel_2d.f90:36-204): z pass on the own sticks, stick -> plane
        // exchange, then y pass and |psi(r)|^2 accumulation on the own planes (exchange slot 0, the main stream)
        const DistHost& D = c->DH;
        const bool p2p = c->P2P.ok && D.nproc > 1;
        if (D.nproc > 1 && !c->nccl_rg && !p2p)
            QEB_FAIL(c, 53, "qeb_sum_band_rho: no NCCL communicator (qeb_nccl_init) or peer mapping (qeb_p2p_import) for the R&G group");
        cudaStream_t st = c->stream;
        cplx* S = (cplx*)c->sbuf_d.p;
        cplx* R = (cplx*)c->rbuf_d.p;
        VlocTables tz = c->T;
        tz.g1 = S;
        tz.g1_zs = (size_t)nb * tz.g1_box;
        VlocTables tp = c->Tp;
        const char* msg = "";
        if (p2p && D.nproc <= 16) {  // fused forward leg, as in run_vloc
            tz.peer_on = 1;
            tz.peer_nb = nb;
            tz.peer_zown = c->fx_zown;
            tz.peer_zi0 = c->fx_zi0;
            for (int q = 0; q < D.nproc; ++q)
                tz.peer_base[q] = c->P2P.peer_R[q] + (size_t)nb * D.nr3p[q] * D.prefix[D.me];
            QEB_CUDA(c, vloc_launch_zpass(true, tz, io, c->cfg, nb, st));
            if ((rc = dist_p2p_sync(D, c->P2P, true, 0, st, &msg))) QEB_FAIL(c, rc, std::string("fft_scatter: ") + msg);
            tp.g1 = R;
            tp.r_on = 1;
            tp.r_nb = nb;
            tp.r_pref = c->fx_rpref;
            tp.r_nsw = c->fx_rnsw;
            tp.r_sidx = c->DD.sidx;
            QEB_CUDA(c, vloc_launch_density(tp, io, c->cfgp, nb, (const double*)w_d, r, st, true));
            // no data flows back: the backward flag words only tell the peers that this rank has consumed its R slice
            if ((rc = dist_p2p_sync(D, c->P2P, false, 0, st, &msg))) QEB_FAIL(c, rc, std::string("fft_scatter: ") + msg);
        } else {
            QEB_CUDA(c, vloc_launch_zpass(true, tz, io, c->cfg, nb, st));
            if ((rc = dist_exchange(D, c->nccl_rg, true, S, R, nb, st, &msg))) QEB_FAIL(c, rc, std::string("fft_scatter: ") + msg);
            QEB_CUDA(c, dist_unpack(c->DD, R, tp.g1, nb, st));
            QEB_CUDA(c, vloc_launch_density(tp, io, c->cfgp, nb, (const double*)w_d, r, st, true));
        }
    }
    if (!rho_dev) {
        QEB_CUDA(c, cudaMemcpyAsync(rho, r, nnr * 8, cudaMemcpyDeviceToHost, c->stream));
        QEB_CUDA(c, cudaStreamSynchronize(c->stream));
        return 0;
    }
    if (!is_device_ptr(evc) || !is_device_ptr(w)) QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    return finish(c);
}

// ---- block-Davidson building blocks ---------------------------
int qeb_set_g_psi_diag(qeb_ctx* c, int n, const double* h_diag, const double* s_diag)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (n < 1 || !h_diag || !s_diag) QEB_FAIL(c, 90, "qeb_set_g_psi_diag: bad arguments");
    QEB_CUDA(c, c->gpsi_d.ensure((size_t)n * 16));
    QEB_CUDA(c, cudaMemcpyAsync(c->gpsi_d.p, h_diag, (size_t)n * 8,
                                is_device_ptr(h_diag) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
    QEB_CUDA(c, cudaMemcpyAsync((double*)c->gpsi_d.p + n, s_diag, (size_t)n * 8,
                                is_device_ptr(s_diag) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
    QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    c->gpsi_n = n;
    return 0;
}

int qeb_g_psi(qeb_ctx* c, int lda, int n, int m, int npol, void* psi, const double* e)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (npol != 1) QEB_FAIL(c, 91, "qeb_g_psi: npol = 2 (noncollinear) is routed to the reference routine by the shim");
    if (n < 1 || lda < n || m < 0 || !psi || !e) QEB_FAIL(c, 91, "qeb_g_psi: bad arguments");
    if (c->gpsi_n < n) QEB_FAIL(c, 92, "qeb_g_psi: qeb_set_g_psi_diag has not been called for this n");
    if (!is_device_ptr(psi)) QEB_FAIL(c, 93, "qeb_g_psi: psi must be a device pointer (the Davidson work arrays live on the device)");
    if (m == 0) return 0;
    const void* ed;
    int rc = stage_in(c, e, (size_t)m * 8, c->dav_small_d, &ed);
    if (rc) return rc;
    QEB_CUDA(c, launch_g_psi(lda, n, m, (cplx*)psi, (const double*)ed, (const double*)c->gpsi_d.p,
                             (const double*)c->gpsi_d.p + c->gpsi_n, c->stream));
    if (!is_device_ptr(e)) QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    return finish(c);
}

int qeb_col_dot(qeb_ctx* c, int n, int m, const void* a, int lda, const void* b, int ldb, double* out)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (n < 1 || m < 0 || lda < n || ldb < n || !a || !b || !out) QEB_FAIL(c, 94, "qeb_col_dot: bad arguments");
    if (!is_device_ptr(a) || !is_device_ptr(b)) QEB_FAIL(c, 93, "qeb_col_dot: a and b must be device pointers");
    if (m == 0) return 0;
    const bool od = is_device_ptr(out);
    double* o = out;
    if (!od) {
        QEB_CUDA(c, c->dav_small_d.ensure((size_t)m * 8));
        o = (double*)c->dav_small_d.p;
    }
    QEB_CUDA(c, launch_col_dot(n, m, (const cplx*)a, lda, (const cplx*)b, ldb, c->gamma ? 1 : 0, c->gstart, o, c->stream));
    if (!od) {
        QEB_CUDA(c, cudaMemcpyAsync(out, o, (size_t)m * 8, cudaMemcpyDeviceToHost, c->stream));
        QEB_CUDA(c, cudaStreamSynchronize(c->stream));
        return 0;
    }
    return finish(c);
}

// C(ka,mb) = A(1:n,1:ka)^H B(1:n,1:mb): the reduced matrices hc/sc of cegterg (ZGEMM 'C','N', cegterg.f90:209-229,
// 478-499) and hr/sr of regterg (2*DGEMM('T','N') - DGER at G=0, regterg.f90:203-221, 402-429) on the calbec kernel.
int qeb_dav_gemm_cn(qeb_ctx* c, int n, const void* A, int lda, int ka, const void* B, int ldb, int mb, void* C, int ldc)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (!c->have_grid) QEB_FAIL(c, 95, "qeb_dav_gemm_cn: qeb_set_grid has not been called (gamma_only / gstart are context state)");
    if (n < 1 || lda < n || ldb < n || ka < 0 || mb < 0 || ldc < ka || !A || !B || !C) QEB_FAIL(c, 95, "qeb_dav_gemm_cn: bad arguments");
    if (!is_device_ptr(A) || !is_device_ptr(B)) QEB_FAIL(c, 93, "qeb_dav_gemm_cn: A and B must be device pointers");
    if (ka == 0 || mb == 0) return 0;
    const size_t el = c->gamma ? 8 : 16;
    int nsplit, gchunk;
    calbec_plan(n, ka, mb, !c->gamma, c->num_sms, &nsplit, &gchunk);
    QEB_CUDA(c, c->part_d.ensure(calbec_workspace_bytes(ka, mb, nsplit, !c->gamma)));
    QEB_CUDA(c, c->misc_stage.ensure((size_t)ka * mb * el));
    QEB_CUDA(c, launch_calbec((const cplx*)A, lda, (const cplx*)B, ldb, n, ka, mb, !c->gamma, c->gstart, nsplit, gchunk,
                              c->part_d.p, c->misc_stage.p, c->stream));
    QEB_CUDA(c, cudaMemcpy2DAsync(C, (size_t)ldc * el, c->misc_stage.p, (size_t)ka * el, (size_t)ka * el, mb,
                                  is_device_ptr(C) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, c->stream));
    if (!is_device_ptr(C)) {
        QEB_CUDA(c, cudaStreamSynchronize(c->stream));
        return 0;
    }
    return finish(c);
}

// C(1:n,1:m) (+)= A(1:n,1:k) B(1:k,1:m): the basis updates of cegterg (ZGEMM 'N','N', cegterg.f90:353-395, 588-632)
// and regterg (DGEMM on the 2*npwx real view, B real) on the add_vuspsi kernel. accumulate == 0: C is overwritten
// (rows n+1..ldc of C are left untouched, as ZGEMM does).
int qeb_dav_gemm_nn(qeb_ctx* c, int n, const void* A, int lda, int k, const void* B, int ldb, int m, void* C, int ldc,
                    int accumulate)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (!c->have_grid) QEB_FAIL(c, 96, "qeb_dav_gemm_nn: qeb_set_grid has not been called");
    if (n < 1 || lda < n || ldc < n || k < 0 || m < 0 || ldb < k || !A || !B || !C) QEB_FAIL(c, 96, "qeb_dav_gemm_nn: bad arguments");
    if (!is_device_ptr(A) || !is_device_ptr(C)) QEB_FAIL(c, 93, "qeb_dav_gemm_nn: A and C must be device pointers");
    if (m == 0) return 0;
    if (!accumulate) QEB_CUDA(c, cudaMemset2DAsync(C, (size_t)ldc * 16, 0, (size_t)n * 16, m, c->stream));
    if (k == 0) return finish(c);
    const size_t el = c->gamma ? 8 : 16;
    const int ldps = (k + 1) & ~1;
    QEB_CUDA(c, c->misc_stage.ensure((size_t)ldps * m * el));
    QEB_CUDA(c, cudaMemsetAsync(c->misc_stage.p, 0, (size_t)ldps * m * el, c->stream));
    QEB_CUDA(c, cudaMemcpy2DAsync(c->misc_stage.p, (size_t)ldps * el, B, (size_t)ldb * el, (size_t)k * el, m,
                                  is_device_ptr(B) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
    QEB_CUDA(c, launch_addvus((const cplx*)A, lda, c->misc_stage.p, ldps, n, k, m, !c->gamma, (cplx*)C, ldc, c->stream));
    if (!is_device_ptr(B)) QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    return finish(c);
}

// ---- peer-to-peer stick <-> plane exchange over NVLink (cudaIpc) ------------
struct P2PBlob {  // what one rank publishes to the others (fits QEB_P2P_BLOB_BYTES)
    cudaIpcMemHandle_t s, r, f;
    unsigned long long s_bytes, r_bytes;
    int batch, nslots, rank, nproc;
};

int qeb_p2p_export(qeb_ctx* c, void* blob)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (!c->dist || !c->have_k) QEB_FAIL(c, 98, "qeb_p2p_export: call after qeb_set_grid_dist + qeb_set_kpoint (the buffers must exist)");
    if (!blob) QEB_FAIL(c, 98, "qeb_p2p_export: null blob");
    static_assert(sizeof(P2PBlob) <= 256, "QEB_P2P_BLOB_BYTES");
    const DistHost& D = c->DH;
    const int ns = c->vloc_nstreams;
    QEB_CUDA(c, c->p2p_flags_d.ensure((size_t)2 * ns * D.nproc * 4 + 64));
    QEB_CUDA(c, cudaMemsetAsync(c->p2p_flags_d.p, 0, c->p2p_flags_d.cap, c->stream));
    QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    P2PBlob b;
    memset(&b, 0, sizeof(b));
    QEB_CUDA(c, cudaIpcGetMemHandle(&b.s, c->sbuf_d.p));
    QEB_CUDA(c, cudaIpcGetMemHandle(&b.r, c->rbuf_d.p));
    QEB_CUDA(c, cudaIpcGetMemHandle(&b.f, c->p2p_flags_d.p));
    b.s_bytes = c->sbuf_d.cap;
    b.r_bytes = c->rbuf_d.cap;
    b.batch = c->batch;
    b.nslots = ns;
    b.rank = D.me;
    b.nproc = D.nproc;
    memset(blob, 0, 256);
    memcpy(blob, &b, sizeof(b));
    c->p2p_exported = true;
    c->P2P = P2PState();
    return 0;
}

int qeb_p2p_import(qeb_ctx* c, const void* blobs, int nranks)
{
    if (!c) return 1;
    ++c->state_epoch;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (!c->p2p_exported) QEB_FAIL(c, 99, "qeb_p2p_import: qeb_p2p_export first");
    const DistHost& D = c->DH;
    if (!blobs || nranks != D.nproc) QEB_FAIL(c, 99, "qeb_p2p_import: one blob per rank of the R&G group expected");
    P2PState& P = c->P2P;
    P = P2PState();
    for (void* m : c->p2p_opened) cudaIpcCloseMemHandle(m);  // mappings of a previous import
    c->p2p_opened.clear();
    P.nslots = c->vloc_nstreams;
    P.batch = c->batch;
    P.peer_S.assign(nranks, nullptr);
    P.peer_R.assign(nranks, nullptr);
    P.peer_flags.assign(nranks, nullptr);
    for (int q = 0; q < nranks; ++q) {
        P2PBlob b;
        memcpy(&b, (const char*)blobs + (size_t)q * 256, sizeof(b));
        if (b.rank != q || b.nproc != nranks) QEB_FAIL(c, 99, "qeb_p2p_import: blobs are not ordered by rank");
        if (b.batch != c->batch || b.nslots != P.nslots) QEB_FAIL(c, 99, "qeb_p2p_import: ranks disagree on batch / streams");
        if (q == D.me) {
            P.peer_S[q] = (cplx*)c->sbuf_d.p;
            P.peer_R[q] = (cplx*)c->rbuf_d.p;
            P.peer_flags[q] = (unsigned*)c->p2p_flags_d.p;
            continue;
        }
        void *ps = nullptr, *pr = nullptr, *pf = nullptr;
        QEB_CUDA(c, cudaIpcOpenMemHandle(&ps, b.s, cudaIpcMemLazyEnablePeerAccess));
        c->p2p_opened.push_back(ps);
        QEB_CUDA(c, cudaIpcOpenMemHandle(&pr, b.r, cudaIpcMemLazyEnablePeerAccess));
        c->p2p_opened.push_back(pr);
        QEB_CUDA(c, cudaIpcOpenMemHandle(&pf, b.f, cudaIpcMemLazyEnablePeerAccess));
        c->p2p_opened.push_back(pf);
        P.peer_S[q] = (cplx*)ps;
        P.peer_R[q] = (cplx*)pr;
        P.peer_flags[q] = (unsigned*)pf;
    }
    QEB_CUDA(c, c->p2p_ptrs_d.ensure((size_t)nranks * sizeof(void*)));
    QEB_CUDA(c, cudaMemcpyAsync(c->p2p_ptrs_d.p, P.peer_flags.data(), (size_t)nranks * sizeof(void*), cudaMemcpyHostToDevice,
                                c->stream));
    QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    P.peer_flags_dev = (unsigned**)c->p2p_ptrs_d.p;
    const char* e = getenv("QEB_P2P");
    P.ok = !(e && atoi(e) == 0);
    return 0;
}

// ---- internal pieces ----------------------------------------------------
static int run_vloc(qeb_ctx* c, int lda, int n, int m, const cplx* psi, cplx* hpsi, bool kinetic)
{
    if (!c->have_v) QEB_FAIL(c, 50, "V_loc: no potential set (qeb_set_vloc)");
    VlocIO io;
    io.psi = psi;
    io.hpsi = hpsi;
    if (kinetic && !c->have_g2kin) QEB_FAIL(c, 55, "h_psi: no g2kin in the context (qeb_set_kpoint or qeb_g2_kin)");
    io.g2kin = kinetic ? (const double*)c->g2kin_d.p : nullptr;
    io.lda = lda;
    io.n = n;
    io.nbands = m;
    io.gamma = c->gamma ? 1 : 0;
    const int nboxes = c->gamma ? (m + 1) / 2 : m;
    if (c->dist) {
        // cft_1z | fft_scatter | cft_2xy * V * cft_2xy | fft_scatter | cft_1z  (fft_parallel_2d.f90:36-204)
        const DistHost& D = c->DH;
        if (D.nproc > 1 && !c->nccl_rg && !c->P2P.ok)
            QEB_FAIL(c, 53, "V_loc: no NCCL communicator (qeb_nccl_init) or peer mapping (qeb_p2p_import) for the R&G group");
        // batches alternate between the V_loc streams (own S/R/G1/G2 slices): the exchange of one batch overlaps
        // the FFT passes of the other. Every rank issues the collectives in the same order.
        const int nbatches_d = (nboxes + c->batch - 1) / c->batch;
        const int nsd = std::min(c->vloc_nstreams, nbatches_d);
        if (nsd > 1) {
            QEB_CUDA(c, cudaEventRecord(c->vev[0], c->stream));
            for (int s = 1; s < nsd; ++s) QEB_CUDA(c, cudaStreamWaitEvent(c->vstream[s], c->vev[0], 0));
        }
        const size_t s_slice = (size_t)D.nsw[D.me] * c->nr3 * c->batch;
        const size_t r_slice = c->Tp.g1_box * c->batch;
        const size_t g2_slice = c->Tp.g2_box * c->batch;
        int ib = 0;
        for (int b0 = 0; b0 < nboxes; b0 += c->batch, ++ib) {
            const int nb = std::min(c->batch, nboxes - b0);
            const int si = ib % nsd;
            cudaStream_t st = si == 0 ? c->stream : c->vstream[si];
            cplx* S = (cplx*)c->sbuf_d.p + si * s_slice;
            cplx* R = (cplx*)c->rbuf_d.p + si * r_slice;
            io.band0 = c->gamma ? 2 * b0 : b0;
            VlocTables tz = c->T;
            tz.g1 = S;
            tz.g1_zs = (size_t)nb * tz.g1_box;
            VlocTables tp = c->Tp;
            tp.g1 += si * r_slice;
            tp.g2 += si * g2_slice;
            const char* msg = "";
            int rc;
            const bool p2p = c->P2P.ok && D.nproc > 1;
            // Fused exchange (peer-to-peer mapping, three-kernel plane path): the z-inverse pass stores straight into the
            // plane owners' R buffers over NVLink, the y passes read and write R itself - no S staging and no repack
            // kernels on the forward leg, no pack kernel on the backward leg (fft_scatter_2d.f90:123,226 folded into the
            // neighbouring passes). QEB_FUSED_X=0: the unfused sequence.
            const bool fused = p2p && !vloc_uses_plane_kernel(tp, c->cfgp) && D.nproc <= 16 && env_flag("QEB_FUSED_X", 1);
            if (fused) {
                tz.peer_on = 1;
                tz.peer_nb = nb;
                tz.peer_zown = c->fx_zown;
                tz.peer_zi0 = c->fx_zi0;
                for (int q = 0; q < D.nproc; ++q)
                    tz.peer_base[q] = c->P2P.peer_R[q] + (size_t)si * ((size_t)D.nst_tot * D.nr3p[q] * c->P2P.batch) +
                                      (size_t)nb * D.nr3p[q] * D.prefix[D.me];
                QEB_CUDA(c, vloc_launch_zpass(true, tz, io, c->cfg, nb, st));
                tz.peer_on = 0;
                if ((rc = dist_p2p_sync(D, c->P2P, true, si, st, &msg))) QEB_FAIL(c, rc, std::string("fft_scatter: ") + msg);
                tp.g1 = R;
                tp.r_on = 1;
                tp.r_nb = nb;
                tp.r_pref = c->fx_rpref;
                tp.r_nsw = c->fx_rnsw;
                tp.r_sidx = c->DD.sidx;
                QEB_CUDA(c, vloc_launch_planes(tp, c->cfgp, nb, st));
                if ((rc = dist_exchange_p2p(D, c->P2P, false, si, S, R, nb, c->nr3, st, &msg)))
                    QEB_FAIL(c, rc, std::string("fft_scatter: ") + msg);
                QEB_CUDA(c, vloc_launch_zpass(false, tz, io, c->cfg, nb, st));
                continue;
            }
            QEB_CUDA(c, vloc_launch_zpass(true, tz, io, c->cfg, nb, st));
            if ((rc = p2p ? dist_exchange_p2p(D, c->P2P, true, si, S, R, nb, c->nr3, st, &msg)
                          : dist_exchange(D, c->nccl_rg, true, S, R, nb, st, &msg)))
                QEB_FAIL(c, rc, std::string("fft_scatter: ") + msg);
            QEB_CUDA(c, dist_unpack(c->DD, R, tp.g1, nb, st));
            QEB_CUDA(c, vloc_launch_planes(tp, c->cfgp, nb, st));
            QEB_CUDA(c, dist_pack(c->DD, tp.g1, R, nb, st));
            if ((rc = p2p ? dist_exchange_p2p(D, c->P2P, false, si, S, R, nb, c->nr3, st, &msg)
                          : dist_exchange(D, c->nccl_rg, false, S, R, nb, st, &msg)))
                QEB_FAIL(c, rc, std::string("fft_scatter: ") + msg);
            QEB_CUDA(c, vloc_launch_zpass(false, tz, io, c->cfg, nb, st));
        }
        for (int s = 1; s < nsd; ++s) {
            QEB_CUDA(c, cudaEventRecord(c->vev[s], c->vstream[s]));
            QEB_CUDA(c, cudaStreamWaitEvent(c->stream, c->vev[s], 0));
        }
        return 0;
    }
    const int nbatches = (nboxes + c->batch - 1) / c->batch;
    const int ns = std::min(c->vloc_nstreams, nbatches);
    if (ns > 1) {  // fork: the side streams start after everything queued so far on the main stream
        QEB_CUDA(c, cudaEventRecord(c->vev[0], c->stream));
        for (int s = 1; s < ns; ++s) QEB_CUDA(c, cudaStreamWaitEvent(c->vstream[s], c->vev[0], 0));
    }
    int ib = 0;
    for (int b0 = 0; b0 < nboxes; b0 += c->batch, ++ib) {
        const int nb = std::min(c->batch, nboxes - b0);
        const int s = ib % ns;
        io.band0 = c->gamma ? 2 * b0 : b0;
        VlocTables T = c->T;
        T.g1 += (size_t)s * c->batch * T.g1_box;
        T.g2 += (size_t)s * c->batch * T.g2_box;
        QEB_CUDA(c, vloc_launch_batch(T, io, c->cfg, nb, s == 0 ? c->stream : c->vstream[s]));
    }
    for (int s = 1; s < ns; ++s) {  // join
        QEB_CUDA(c, cudaEventRecord(c->vev[s], c->vstream[s]));
        QEB_CUDA(c, cudaStreamWaitEvent(c->stream, c->vev[s], 0));
    }
    return 0;
}

static int ensure_becp(qeb_ctx* c, int nkb, int m)
{
    const size_t el = c->gamma ? 8 : 16;
    QEB_CUDA(c, c->becp_d.ensure((size_t)nkb * m * el));
    const int ldps = (nkb + 1) & ~1;
    const size_t need = (size_t)ldps * m * el;
    if (need > c->ps_d.cap) {
        QEB_CUDA(c, c->ps_d.ensure(need));
    }
    QEB_CUDA(c, cudaMemsetAsync(c->ps_d.p, 0, need, c->stream));
    return 0;
}

static int run_calbec(qeb_ctx* c, const cplx* beta, int npwx, int nkb, const cplx* psi, int lda, int n, int m, void* becp_out)
{
    int nsplit, gchunk;
    calbec_plan(n, nkb, m, !c->gamma, c->num_sms, &nsplit, &gchunk);
    QEB_CUDA(c, c->part_d.ensure(calbec_workspace_bytes(nkb, m, nsplit, !c->gamma)));
    QEB_CUDA(c, launch_calbec(beta, npwx, psi, lda, n, nkb, m, !c->gamma, c->gstart, nsplit, gchunk, c->part_d.p,
                              becp_out, c->stream));
    if (c->dist && c->DH.nproc > 1) {
        // mp_sum(becp, intra_bgrp_comm): Modules/becmod.f90:395-399 (k), :327-331 (Gamma)
        const NcclApi& A = nccl_api();
        if (!A.ok || !c->nccl_rg) QEB_FAIL(c, 54, "calbec: no NCCL communicator for the R&G group (qeb_nccl_init)");
        const int rc = A.AllReduce(becp_out, becp_out, (size_t)nkb * m * (c->gamma ? 1 : 2), kNcclDouble, kNcclSum,
                                   c->nccl_rg, c->stream);
        if (rc) QEB_FAIL(c, 900 + rc, std::string("ncclAllReduce(becp): ") + A.GetErrorString(rc));
    }
    return 0;
}

static int run_addvus(qeb_ctx* c, int lda, int n, int m, cplx* hpsi, int col0 = 0, const double* dmat = nullptr)
{
    if (!c->have_nl) QEB_FAIL(c, 51, "add_vuspsi: qeb_set_nonlocal has not been called");
    if (!dmat) dmat = (const double*)c->deeq_d.p;
    const int nkb = c->nkb;
    if (nkb > c->nl_rows) QEB_FAIL(c, 52, "add_vuspsi: vkb has projectors not covered by ofsbeta/nh of qeb_set_nonlocal");
    const int ldps = (nkb + 1) & ~1;
    const size_t el = c->gamma ? 8 : 16;
    const char* becp = (const char*)c->becp_d.p + (size_t)col0 * nkb * el;
    char* ps = (char*)c->ps_d.p + (size_t)col0 * ldps * el;
    QEB_CUDA(c, launch_apply_deeq(becp, nkb, m, !c->gamma, c->row_atom_d, c->ofs_d, c->nh_d,
                                  dmat, c->nhm, ps, ldps, c->stream));
    QEB_CUDA(c, launch_addvus(c->vkb_d, c->npwx, ps, ldps, n, nkb, m, !c->gamma, hpsi, lda, c->stream));
    return 0;
}

static int check_call(qeb_ctx* c, int lda, int n, int m, const void* psi, const void* hpsi, const char* who)
{
    if (!c->have_k) QEB_FAIL(c, 60, std::string(who) + ": qeb_set_kpoint has not been called");
    if (n != c->npw) QEB_FAIL(c, 61, std::string(who) + ": n differs from the npw of qeb_set_kpoint");
    if (lda < n || m < 0 || !psi || !hpsi) QEB_FAIL(c, 62, std::string(who) + ": bad arguments");
    return 0;
}

// columns [col0, col0+m) of a call over m_total columns; psi_d / hpsi_d point at column col0. The projections of
// the whole call live in becp_d (column j at j*nkb): the caller runs ensure_becp(m_total) once when col0 >= 0 is
// used with m_total > m.
static int h_psi_cols(qeb_ctx* c, int lda, int n, int m, const cplx* psi_d, cplx* hpsi_d, bool do_vloc, bool do_nl,
                      bool kinetic, int col0 = 0, int m_total = -1)
{
    int rc;
    if (kinetic && lda > n) QEB_CUDA(c, launch_zero_pad_rows(hpsi_d, lda, n, m, c->stream));
    if (do_vloc && (rc = run_vloc(c, lda, n, m, psi_d, hpsi_d, kinetic))) return rc;
    if (do_nl && c->nkb > 0 && !c->vkb_d) QEB_FAIL(c, 56, "h_psi: no vkb in the context (qeb_set_kpoint or qeb_init_us_2)");
    if (do_nl && c->nkb > 0) {
        if (m_total < 0 && (rc = ensure_becp(c, c->nkb, m))) return rc;
        char* becp = (char*)c->becp_d.p + (size_t)col0 * c->nkb * (c->gamma ? 8 : 16);
        if ((rc = run_calbec(c, c->vkb_d, c->npwx, c->nkb, psi_d, lda, n, m, becp))) return rc;
        c->becp_m = m_total < 0 ? m : (col0 + m);
        if ((rc = run_addvus(c, lda, n, m, hpsi_d, col0))) return rc;
    }
    if (kinetic && c->gamma && c->gstart == 2) QEB_CUDA(c, launch_zero_im_g0(hpsi_d, lda, m, c->stream));
    return 0;
}

// Page-locking of caller memory is EXPLICIT (qeb_pin_host / qeb_unpin_host, bracketing the lifetime of the caller's
// work arrays): the library never registers memory on its own, because it cannot see the caller free and re-allocate
// an array at the same address. Unpinned (pageable) arrays go through the same pipeline; cudaMemcpyAsync then stages
// them through the driver's bounce buffers and the copies overlap less.

// h_psi with HOST psi and hpsi: column chunks flow through three streams, so the H2D copy of chunk i+1 and the D2H
// copy of chunk i-1 overlap the kernels of chunk i. Every column is independent in h_psi (h_psi.f90:53-67 splits
// columns the same way over band groups), so the result is the one of the single call.
static int h_psi_host_pipeline(qeb_ctx* c, int lda, int n, int m, const void* psi, void* hpsi)
{
    int rc;
    const size_t col = (size_t)lda * 16;
    int mc = c->pipeline_chunk;
    if (mc <= 0) {
        // whole band tiles of BOTH projector kernels (calbec: 48 bands, add_vuspsi: 64): multiples of 192; measured at
        // AUSURF112 m=800 (tools/e2e_sweep.py): 96 -> 101.0, 128 -> 95.7, 192 -> 93.1, 256 -> 100.0, 400 -> 112.9 ms;
        // with the sub-chunked first copy-in / last copy-outs below: 192 -> 87.9, 256 -> 91.5, 272 -> 93.0, 400 -> 91.2 ms
        mc = m >= 384 ? 192 : ((m + 7) / 8 + 63) / 64 * 64;
        if (mc < 64) mc = 64;
    }
    if (mc > m) mc = m;
    if (mc & 1) ++mc;  // Gamma: band pairs stay together
    // chunk schedule: the first copy-in and the last copy-out are the only transfers nothing overlaps, so in the
    // automatic schedule the first and the last chunk are short (64 columns = one GEMM band tile)
    std::vector<int> sizes;
    {
        int rem = m;
        // (measured: no gain at AUSURF112, 96.9 vs 95.7 ms, because the short chunks run the GEMMs less efficiently;
        // kept behind QEB_PIPE_RAMP for hosts with slower links)
        const bool ramp = getenv("QEB_PIPE_RAMP") && c->pipeline_chunk <= 0 && m >= 4 * 64 && mc > 64;
        if (ramp) {
            sizes.push_back(64);
            rem -= 128;  // first + last
        }
        while (rem > 0) {
            const int t = std::min(mc, rem);
            sizes.push_back(t);
            rem -= t;
        }
        if (ramp) sizes.push_back(64);
    }
    const int nch = (int)sizes.size();
    if (!c->s_in) QEB_CUDA(c, cudaStreamCreateWithFlags(&c->s_in, cudaStreamNonBlocking));
    if (!c->s_out) QEB_CUDA(c, cudaStreamCreateWithFlags(&c->s_out, cudaStreamNonBlocking));
    while ((int)c->ev.size() < 2 * nch) {
        cudaEvent_t e;
        QEB_CUDA(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        c->ev.push_back(e);
    }
    QEB_CUDA(c, c->psi_stage.ensure(col * m));
    QEB_CUDA(c, c->hpsi_stage.ensure(col * m));
    if (c->nkb > 0 && (rc = ensure_becp(c, c->nkb, m))) return rc;
    // the staging buffers may still be read by work queued earlier on the compute stream (async callers)
    QEB_CUDA(c, cudaEventRecord(c->ev[0], c->stream));
    QEB_CUDA(c, cudaStreamWaitEvent(c->s_in, c->ev[0], 0));
    const char* ph = (const char*)psi;
    char* hh = (char*)hpsi;
    char* pd = (char*)c->psi_stage.p;
    char* hd = (char*)c->hpsi_stage.p;
    int c0 = 0;
    // The copy-in of the first chunk is the one transfer no compute can hide. Its V_loc part only needs the columns
    // that have landed, so the first chunk is copied in sub-chunks of `sub` columns and the V_loc passes follow them
    // one sub-chunk behind; the projector GEMMs then run on the whole chunk (serial descriptor only: the distributed
    // path keeps one collective schedule per chunk).
    int sub = 40;
    if (const char* e = getenv("QEB_PIPE_SUB")) sub = atoi(e);
    if (sub & 1) ++sub;
    int tsub = 64;  // one band tile of the add_vuspsi kernels
    if (const char* e = getenv("QEB_PIPE_TAIL")) tsub = atoi(e);
    if (tsub & 1) ++tsub;
    for (int i = 0; i < nch; c0 += sizes[i], ++i) {
        const int mm = sizes[i];
        if (i == 0 && sub > 0 && mm > sub && !c->dist) {
            while ((int)c->ev_sub.size() < (mm + sub - 1) / sub) {
                cudaEvent_t e;
                QEB_CUDA(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
                c->ev_sub.push_back(e);
            }
            for (int s0 = 0, k = 0; s0 < mm; s0 += sub, ++k) {
                const int ms = std::min(sub, mm - s0);
                QEB_CUDA(c, cudaMemcpyAsync(pd + col * s0, ph + col * s0, col * ms, cudaMemcpyHostToDevice, c->s_in));
                QEB_CUDA(c, cudaEventRecord(c->ev_sub[k], c->s_in));
                QEB_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_sub[k], 0));
                if ((rc = h_psi_cols(c, lda, n, ms, (const cplx*)(pd + col * s0), (cplx*)(hd + col * s0), true, false, true)))
                    return rc;
            }
            if ((rc = h_psi_cols(c, lda, n, mm, (const cplx*)pd, (cplx*)hd, false, true, true, 0, m))) return rc;
            QEB_CUDA(c, cudaEventRecord(c->ev[1], c->stream));
            QEB_CUDA(c, cudaStreamWaitEvent(c->s_out, c->ev[1], 0));
            QEB_CUDA(c, cudaMemcpyAsync(hh, hd, col * mm, cudaMemcpyDeviceToHost, c->s_out));
            continue;
        }
        QEB_CUDA(c, cudaMemcpyAsync(pd + col * c0, ph + col * c0, col * mm, cudaMemcpyHostToDevice, c->s_in));
        QEB_CUDA(c, cudaEventRecord(c->ev[2 * i], c->s_in));
        QEB_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev[2 * i], 0));
        if (i > 0 && i >= nch - 2 && tsub > 0 && mm > tsub && !c->dist && c->nkb > 0 && c->vkb_d && c->have_nl) {
            // The copy-out of the last chunks is the other transfer little compute is left to hide: add_vuspsi, the last
            // kernel of a chunk, runs in sub-chunks of one band tile and each sub-chunk is copied out behind it.
            const cplx* pc = (const cplx*)(pd + col * c0);
            cplx* hc = (cplx*)(hd + col * c0);
            if ((rc = h_psi_cols(c, lda, n, mm, pc, hc, true, false, true))) return rc;
            char* becp = (char*)c->becp_d.p + (size_t)c0 * c->nkb * (c->gamma ? 8 : 16);
            if ((rc = run_calbec(c, c->vkb_d, c->npwx, c->nkb, pc, lda, n, mm, becp))) return rc;
            c->becp_m = c0 + mm;
            while ((int)c->ev_tail.size() < (mm + tsub - 1) / tsub) {
                cudaEvent_t e;
                QEB_CUDA(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
                c->ev_tail.push_back(e);
            }
            for (int s0 = 0, k = 0; s0 < mm; s0 += tsub, ++k) {
                const int ms = std::min(tsub, mm - s0);
                cplx* hs = hc + (size_t)s0 * lda;
                if ((rc = run_addvus(c, lda, n, ms, hs, c0 + s0))) return rc;
                if (c->gamma && c->gstart == 2) QEB_CUDA(c, launch_zero_im_g0(hs, lda, ms, c->stream));
                QEB_CUDA(c, cudaEventRecord(c->ev_tail[k], c->stream));
                QEB_CUDA(c, cudaStreamWaitEvent(c->s_out, c->ev_tail[k], 0));
                QEB_CUDA(c, cudaMemcpyAsync(hh + col * (c0 + s0), hs, col * ms, cudaMemcpyDeviceToHost, c->s_out));
            }
            continue;
        }
        if ((rc = h_psi_cols(c, lda, n, mm, (const cplx*)(pd + col * c0), (cplx*)(hd + col * c0), true, true, true, c0, m)))
            return rc;
        QEB_CUDA(c, cudaEventRecord(c->ev[2 * i + 1], c->stream));
        QEB_CUDA(c, cudaStreamWaitEvent(c->s_out, c->ev[2 * i + 1], 0));
        QEB_CUDA(c, cudaMemcpyAsync(hh + col * c0, hd + col * c0, col * mm, cudaMemcpyDeviceToHost, c->s_out));
    }
    QEB_CUDA(c, cudaStreamSynchronize(c->s_out));
    QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    return 0;
}

// ---- CUDA-graph replay of h_psi -------------------------------------------------------------------------------
// The small BASELINE configs are launch-bound (Si-2: 8 launches for 3 MB of work, 0.12 ms per call): the third call
// with the same arguments and context state is captured into a graph (the V_loc side streams fork and join inside the
// capture) and later calls replay it with one launch. The key holds everything a launch depends on: the arguments,
// the setters' epoch, and the addresses of the buffers the kernels receive.
static std::vector<unsigned long long> graph_key(qeb_ctx* c, int lda, int n, int m, const void* psi, void* hpsi)
{
    auto P = [](const void* p) { return (unsigned long long)(uintptr_t)p; };
    return {(unsigned long long)lda, (unsigned long long)n, (unsigned long long)m, P(psi), P(hpsi), c->state_epoch,
            P(c->vkb_d), P(c->g2kin_d.p), P(c->vp_d.p), P(c->vpl_d.p), P(c->deeq_d.p), P(c->nlint_d.p), P(c->tab_d.p),
            P(c->plan_d.p), P(c->g1_d.p), P(c->g2_d.p), P(c->becp_d.p), P(c->ps_d.p), P(c->part_d.p),
            (unsigned long long)c->nkb, (unsigned long long)c->batch};
}

static bool graph_wanted(qeb_ctx* c, int n, int m)
{
    const char* e = getenv("QEB_GRAPH");  // 0: off, 2: every device-pointer call, otherwise automatic (read per call)
    c->graph_mode = e ? (atoi(e) == 0 ? 0 : (atoi(e) == 2 ? 1 : 2)) : 2;
    if (c->graph_mode == 0 || c->dist) return false;
    if (c->graph_mode == 1) return true;
    return (long long)n * m <= (1LL << 17);  // automatic: calls whose kernels are shorter than their launches
}

static void graph_clear(qeb_ctx* c)
{
    for (auto& g : c->graphs)
        if (g.exec) cudaGraphExecDestroy(g.exec);
    c->graphs.clear();
}

// returns 0 and sets *done when the call was served by a graph (captured now or replayed)
static int h_psi_graph(qeb_ctx* c, int lda, int n, int m, const cplx* psi_d, cplx* hpsi_d, bool* done)
{
    *done = false;
    const std::vector<unsigned long long> key = graph_key(c, lda, n, m, psi_d, hpsi_d);
    qeb_ctx::GraphEntry* ent = nullptr;
    for (auto& g : c->graphs)
        if (g.key == key) ent = &g;
    if (!ent) {
        if (c->graphs.size() >= 64) graph_clear(c);  // a caller that never repeats a call
        c->graphs.push_back({});
        ent = &c->graphs.back();
        ent->key = key;
    }
    if (ent->exec) {
        QEB_CUDA(c, cudaGraphLaunch(ent->exec, c->stream));
        *done = true;
        return 0;
    }
    if (++ent->seen < 2) return 0;  // the first sight of a key runs eagerly (the very first call also sizes the work
                                    // buffers, which changes the key): capture at the second
    cudaGraph_t graph = nullptr;
    QEB_CUDA(c, cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
    const int rc = h_psi_cols(c, lda, n, m, psi_d, hpsi_d, true, true, true);
    const cudaError_t e = cudaStreamEndCapture(c->stream, &graph);
    if (rc != 0 || e != cudaSuccess || !graph) {
        if (graph) cudaGraphDestroy(graph);
        cudaGetLastError();
        ent->seen = -(1 << 30);  // never try this key again
        return 0;                // the eager path runs (and reports its own error, if any, with a clean stream)
    }
    cudaGraphExec_t exec = nullptr;
    const cudaError_t ei = cudaGraphInstantiate(&exec, graph, 0);
    cudaGraphDestroy(graph);
    if (ei != cudaSuccess) {
        cudaGetLastError();
        ent->seen = -(1 << 30);
        return 0;
    }
    ent->exec = exec;
    QEB_CUDA(c, cudaGraphLaunch(exec, c->stream));
    *done = true;
    return 0;
}

int qeb_h_psi(qeb_ctx* c, int lda, int n, int m, const void* psi, void* hpsi)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    int rc = check_call(c, lda, n, m, psi, hpsi, "qeb_h_psi");
    if (rc) return rc;
    if (m == 0) return 0;
    const size_t bytes = (size_t)lda * m * 16;
    const bool out_dev = is_device_ptr(hpsi);
    // (distributed grids: the choice must not depend on the rank's own row count - every rank has to issue the same
    // sequence of collectives)
    if (!out_dev && !is_device_ptr(psi) && m >= c->pipeline_min_cols &&
        (bytes >= ((size_t)8 << 20) || c->pipeline_chunk > 0 || c->dist))
        return h_psi_host_pipeline(c, lda, n, m, psi, hpsi);
    const void* psi_d;
    if ((rc = stage_in(c, psi, bytes, c->psi_stage, &psi_d))) return rc;
    cplx* hpsi_d = (cplx*)hpsi;
    if (!out_dev) {
        QEB_CUDA(c, c->hpsi_stage.ensure(bytes));
        hpsi_d = (cplx*)c->hpsi_stage.p;
    }
    bool replayed = false;
    if (out_dev && is_device_ptr(psi) && graph_wanted(c, n, m)) {
        if ((rc = h_psi_graph(c, lda, n, m, (const cplx*)psi_d, hpsi_d, &replayed))) return rc;
        if (replayed) {
            if (c->nkb > 0) c->becp_m = m;
            return finish(c);
        }
    }
    if ((rc = h_psi_cols(c, lda, n, m, (const cplx*)psi_d, hpsi_d, true, true, true))) return rc;
    if (!out_dev) {
        QEB_CUDA(c, cudaMemcpyAsync(hpsi, hpsi_d, bytes, cudaMemcpyDeviceToHost, c->stream));
        QEB_CUDA(c, cudaStreamSynchronize(c->stream));
        return 0;
    }
    return finish(c);
}

// hs_psi(lda,n,m,psi,hpsi,spsi) (PW/src/hs_psi.f90:7-40) and, with m = 1, hs_1psi (PW/src/hs_1psi.f90:7-60): H and S on
// the same block, S re-using the projections H just left in the context.
int qeb_hs_psi(qeb_ctx* c, int lda, int n, int m, const void* psi, void* hpsi, void* spsi)
{
    if (!c) return 1;
    if (!spsi) QEB_FAIL(c, 62, "qeb_hs_psi: bad arguments");
    const int rc = qeb_h_psi(c, lda, n, m, psi, hpsi);
    if (rc) return rc;
    return qeb_s_psi(c, lda, n, m, psi, spsi);
}

// h_psi of several k-points in one call (the k loop of PW/src/c_bands.f90:95-150 for the launch-bound configs): context
// k applies ITS Hamiltonian to psi[k] on its own stream (graph replay when warm), all streams are joined at the end.
// all k-points of a multi call in ONE graph: the capture starts on the first context's stream, every other context's
// stream is forked from it with an event, runs its own h_psi (which forks its V_loc side streams in turn) and joins
// back. One cudaGraphLaunch then replaces nk x ~8 kernel launches and the k-points run as parallel branches.
static int h_psi_multi_graph(int nk, qeb_ctx* const* ctxs, const int* lda, const int* n, const int* m, const void* const* psi,
                             void* const* hpsi, bool* done)
{
    *done = false;
    qeb_ctx* c0 = ctxs[0];
    std::vector<unsigned long long> key;
    for (int k = 0; k < nk; ++k) {
        qeb_ctx* c = ctxs[k];
        if (!c || c->device != c0->device || !c->have_k || n[k] != c->npw || lda[k] < n[k] || m[k] < 1 || !is_device_ptr(psi[k]) ||
            !is_device_ptr(hpsi[k]) || !graph_wanted(c, n[k], m[k]))
            return 0;
        const auto kk = graph_key(c, lda[k], n[k], m[k], psi[k], hpsi[k]);
        key.push_back((unsigned long long)(uintptr_t)c);
        key.insert(key.end(), kk.begin(), kk.end());
    }
    qeb_ctx::GraphEntry* ent = nullptr;
    for (auto& g : c0->mgraphs)
        if (g.key == key) ent = &g;
    if (!ent) {
        if (c0->mgraphs.size() >= 16) {
            for (auto& g : c0->mgraphs)
                if (g.exec) cudaGraphExecDestroy(g.exec);
            c0->mgraphs.clear();
        }
        c0->mgraphs.push_back({});
        ent = &c0->mgraphs.back();
        ent->key = key;
    }
    if (!ent->exec) {
        if (++ent->seen < 2) return 0;  // as in h_psi_graph
        for (int k = 0; k < nk; ++k)
            for (int j = 0; j < 2; ++j)
                if (!ctxs[k]->mev[j]) QEB_CUDA(c0, cudaEventCreateWithFlags(&ctxs[k]->mev[j], cudaEventDisableTiming));
        cudaGraph_t graph = nullptr;
        QEB_CUDA(c0, cudaStreamBeginCapture(c0->stream, cudaStreamCaptureModeThreadLocal));
        int rc = 0;
        cudaError_t e = cudaEventRecord(c0->mev[0], c0->stream);
        for (int k = 0; k < nk && rc == 0 && e == cudaSuccess; ++k) {
            qeb_ctx* c = ctxs[k];
            if (k > 0 && c->stream != c0->stream) e = cudaStreamWaitEvent(c->stream, c0->mev[0], 0);
            if (e == cudaSuccess) rc = h_psi_cols(c, lda[k], n[k], m[k], (const cplx*)psi[k], (cplx*)hpsi[k], true, true, true);
            if (rc == 0 && e == cudaSuccess && k > 0 && c->stream != c0->stream) {
                e = cudaEventRecord(c->mev[1], c->stream);
                if (e == cudaSuccess) e = cudaStreamWaitEvent(c0->stream, c->mev[1], 0);
            }
        }
        const cudaError_t ee = cudaStreamEndCapture(c0->stream, &graph);
        cudaGraphExec_t exec = nullptr;
        if (rc != 0 || e != cudaSuccess || ee != cudaSuccess || !graph ||
            cudaGraphInstantiate(&exec, graph, 0) != cudaSuccess) {
            if (graph) cudaGraphDestroy(graph);
            if (getenv("QEB_GRAPH_DEBUG"))
                fprintf(stderr, "qeb_h_psi_multi: graph capture failed (rc %d: %s; cuda %s / %s), eager launches from now on\n", rc,
                        rc ? ctxs[0]->err.c_str() : "", cudaGetErrorString(e), cudaGetErrorString(ee));
            cudaGetLastError();
            ent->seen = -(1 << 30);  // never again for this key; the eager path serves the call
            return 0;
        }
        cudaGraphDestroy(graph);
        ent->exec = exec;
    }
    QEB_CUDA(c0, cudaGraphLaunch(ent->exec, c0->stream));
    for (int k = 0; k < nk; ++k)
        if (ctxs[k]->nkb > 0) ctxs[k]->becp_m = m[k];
    *done = true;
    return 0;
}

int qeb_h_psi_multi(int nk, qeb_ctx* const* ctxs, const int* lda, const int* n, const int* m, const void* const* psi,
                    void* const* hpsi)
{
    if (nk < 0 || (nk > 0 && (!ctxs || !lda || !n || !m || !psi || !hpsi))) return 1;
    if (nk == 0) return 0;
    for (int k = 0; k < nk; ++k)
        if (!ctxs[k]) return 1;
    int rc = 0;
    bool done = false;
    cudaSetDevice(ctxs[0]->device);
    if ((rc = h_psi_multi_graph(nk, ctxs, lda, n, m, psi, hpsi, &done))) return rc;
    if (done) {  // everything hangs off the first context's stream
        if (!ctxs[0]->async) {
            const cudaError_t e = cudaStreamSynchronize(ctxs[0]->stream);
            if (e != cudaSuccess) {
                ctxs[0]->err = std::string("qeb_h_psi_multi: ") + cudaGetErrorString(e);
                return 100 + (int)e;
            }
        }
        return 0;
    }
    for (int k = 0; k < nk && rc == 0; ++k) {
        const bool was = ctxs[k]->async;
        ctxs[k]->async = true;  // queue only; joined below
        rc = qeb_h_psi(ctxs[k], lda[k], n[k], m[k], psi[k], hpsi[k]);
        ctxs[k]->async = was;
        if (rc && k > 0) ctxs[0]->err = "qeb_h_psi_multi: k-point " + std::to_string(k) + ": " + ctxs[k]->err;
    }
    // join: every other context's stream first, then the first context's unless it is asynchronous
    for (int k = nk - 1; k >= 0; --k) {
        if (k == 0 && ctxs[0]->async) continue;
        cudaSetDevice(ctxs[k]->device);
        const cudaError_t e = cudaStreamSynchronize(ctxs[k]->stream);
        if (e != cudaSuccess && rc == 0) {
            ctxs[0]->err = std::string("qeb_h_psi_multi: ") + cudaGetErrorString(e);
            rc = 100 + (int)e;
        }
    }
    return rc;
}

static int vloc_entry(qeb_ctx* c, int lda, int n, int m, const void* psi, const double* v, void* hpsi, bool want_gamma,
                      const char* who)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    int rc = check_call(c, lda, n, m, psi, hpsi, who);
    if (rc) return rc;
    if (want_gamma != c->gamma) QEB_FAIL(c, 63, std::string(who) + ": grid was set with a different gamma_only");
    if (m == 0) return 0;
    if (v && (rc = prepare_v(c, v))) return rc;
    const size_t bytes = (size_t)lda * m * 16;
    const void* psi_d;
    if ((rc = stage_in(c, psi, bytes, c->psi_stage, &psi_d))) return rc;
    const bool out_dev = is_device_ptr(hpsi);
    const void* hin;
    if ((rc = stage_in(c, hpsi, bytes, c->hpsi_stage, &hin))) return rc;  // INOUT
    cplx* hpsi_d = (cplx*)hin;
    if ((rc = h_psi_cols(c, lda, n, m, (const cplx*)psi_d, hpsi_d, true, false, false))) return rc;
    if (!out_dev) {
        QEB_CUDA(c, cudaMemcpyAsync(hpsi, hpsi_d, bytes, cudaMemcpyDeviceToHost, c->stream));
        QEB_CUDA(c, cudaStreamSynchronize(c->stream));
        return 0;
    }
    return finish(c);
}

// m spinors = 2m component columns of leading dimension lda through the 2x2 potential; kinetic: hpsi = g2kin*psi + ...
static int run_vloc_nc(qeb_ctx* c, int lda, int n, int m, const cplx* psi_d, cplx* hpsi_d, bool kinetic)
{
    if (!c->have_vnc) QEB_FAIL(c, 50, "vloc_psi_nc: no 2x2 potential set (qeb_set_vloc_nc)");
    if (2 * c->cfg.smem_x > c->max_smem) QEB_FAIL(c, 64, "vloc_psi_nc: x tile of this grid does not fit twice in shared memory");
    if (kinetic && !c->have_g2kin) QEB_FAIL(c, 55, "h_psi_nc: no g2kin in the context (qeb_set_kpoint or qeb_g2_kin)");
    VlocIO io;
    io.psi = psi_d;
    io.hpsi = hpsi_d;
    io.g2kin = kinetic ? (const double*)c->g2kin_d.p : nullptr;
    io.lda = lda;
    io.n = n;
    io.nbands = 2 * m;
    io.gamma = 0;
    const int bands_per_batch = std::max(1, c->batch / 2);
    for (int b0 = 0; b0 < m; b0 += bands_per_batch) {
        const int nb = std::min(bands_per_batch, m - b0);
        io.band0 = 2 * b0;
        QEB_CUDA(c, vloc_launch_batch_nc(c->T, io, c->cfg, nb, c->stream));
    }
    return 0;
}

// vloc_psi_nc with domag (PW/src/vloc_psi_acc.f90:269-369): psi(lda*2, m), hpsi(lda, 2, m) INOUT. The two components of a
// band are columns 2b, 2b+1 of leading dimension lda, i.e. boxes 2b, 2b+1 of the batch.
int qeb_vloc_psi_nc(qeb_ctx* c, int lda, int n, int m, const void* psi, const double* v4, void* hpsi)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    int rc = check_call(c, lda, n, m, psi, hpsi, "qeb_vloc_psi_nc");
    if (rc) return rc;
    if (c->dist || c->gamma) QEB_FAIL(c, 63, "qeb_vloc_psi_nc: serial k-point descriptor only");
    if (m == 0) return 0;
    if (v4 && (rc = prepare_v_nc(c, v4))) return rc;
    const size_t bytes = (size_t)lda * 2 * m * 16;
    const void* psi_d;
    if ((rc = stage_in(c, psi, bytes, c->psi_stage, &psi_d))) return rc;
    const bool out_dev = is_device_ptr(hpsi);
    const void* hin;
    if ((rc = stage_in(c, hpsi, bytes, c->hpsi_stage, &hin))) return rc;  // INOUT
    if ((rc = run_vloc_nc(c, lda, n, m, (const cplx*)psi_d, (cplx*)hin, false))) return rc;
    if (!out_dev) {
        QEB_CUDA(c, cudaMemcpyAsync(hpsi, hin, bytes, cudaMemcpyDeviceToHost, c->stream));
        QEB_CUDA(c, cudaStreamSynchronize(c->stream));
        return 0;
    }
    return finish(c);
}

// deeq_nc(nhm,nhm,nat,4) complex of add_vuspsi_nc (PW/src/add_vuspsi.f90:174-234); nat/nhm as in qeb_set_nonlocal
int qeb_set_nonlocal_nc(qeb_ctx* c, const void* deeq_nc)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (!c->have_nl) QEB_FAIL(c, 51, "qeb_set_nonlocal_nc: qeb_set_nonlocal has not been called (nat, nhm, nh, ofsbeta)");
    if (!deeq_nc) QEB_FAIL(c, 40, "qeb_set_nonlocal_nc: null deeq_nc");
    const size_t nd = (size_t)c->nhm * c->nhm * c->nat * 4;
    QEB_CUDA(c, c->dnc_d.ensure(nd * 16 + 16));
    if (nd)
        QEB_CUDA(c, cudaMemcpyAsync(c->dnc_d.p, deeq_nc, nd * 16,
                                    is_device_ptr(deeq_nc) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
    QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    c->have_dnc = true;
    return 0;
}

// h_psi_ for noncollinear spinors with magnetisation / spin-orbit (PW/src/h_psi.f90:138-148,184-190,225-229):
// kinetic term on both components, vloc_psi_nc, calbec_nc (= calbec on 2m columns, becmod.f90:575-616), add_vuspsi_nc.
// psi(lda*2,m) -> hpsi(lda*2,m); leaves becp%nc(nkb,2,m) in the context (qeb_get_becp with 2m columns).
int qeb_h_psi_nc(qeb_ctx* c, int lda, int n, int m, const void* psi, void* hpsi)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    int rc = check_call(c, lda, n, m, psi, hpsi, "qeb_h_psi_nc");
    if (rc) return rc;
    if (c->dist || c->gamma) QEB_FAIL(c, 63, "qeb_h_psi_nc: serial k-point descriptor only");
    if (m == 0) return 0;
    const int nkb = c->nkb;
    if (nkb > 0 && !c->vkb_d) QEB_FAIL(c, 56, "qeb_h_psi_nc: no vkb in the context (qeb_set_kpoint or qeb_init_us_2)");
    if (nkb > 0 && !c->have_dnc) QEB_FAIL(c, 51, "qeb_h_psi_nc: no deeq_nc in the context (qeb_set_nonlocal_nc)");
    if (nkb > c->nl_rows) QEB_FAIL(c, 52, "qeb_h_psi_nc: vkb has projectors not covered by ofsbeta/nh of qeb_set_nonlocal");
    const size_t bytes = (size_t)lda * 2 * m * 16;
    const bool out_dev = is_device_ptr(hpsi);
    const void* psi_d;
    if ((rc = stage_in(c, psi, bytes, c->psi_stage, &psi_d))) return rc;
    cplx* hpsi_d = (cplx*)hpsi;
    if (!out_dev) {
        QEB_CUDA(c, c->hpsi_stage.ensure(bytes));
        hpsi_d = (cplx*)c->hpsi_stage.p;
    }
    if (lda > n) QEB_CUDA(c, launch_zero_pad_rows(hpsi_d, lda, n, 2 * m, c->stream));
    if ((rc = run_vloc_nc(c, lda, n, m, (const cplx*)psi_d, hpsi_d, true))) return rc;
    if (nkb > 0) {
        if ((rc = ensure_becp(c, nkb, 2 * m))) return rc;
        if ((rc = run_calbec(c, c->vkb_d, c->npwx, nkb, (const cplx*)psi_d, lda, n, 2 * m, c->becp_d.p))) return rc;
        c->becp_m = 2 * m;
        const int ldps = (nkb + 1) & ~1;
        QEB_CUDA(c, launch_apply_deeq_nc(c->becp_d.p, nkb, m, c->row_atom_d, c->ofs_d, c->nh_d, c->dnc_d.p, c->nhm, c->nat,
                                         c->ps_d.p, ldps, c->stream));
        QEB_CUDA(c, launch_addvus(c->vkb_d, c->npwx, c->ps_d.p, ldps, n, nkb, 2 * m, true, hpsi_d, lda, c->stream));
    }
    if (!out_dev) {
        QEB_CUDA(c, cudaMemcpyAsync(hpsi, hpsi_d, bytes, cudaMemcpyDeviceToHost, c->stream));
        QEB_CUDA(c, cudaStreamSynchronize(c->stream));
        return 0;
    }
    return finish(c);
}

int qeb_vloc_psi_k(qeb_ctx* c, int lda, int n, int m, const void* psi, const double* v, void* hpsi)
{
    return vloc_entry(c, lda, n, m, psi, v, hpsi, false, "qeb_vloc_psi_k");
}
int qeb_vloc_psi_gamma(qeb_ctx* c, int lda, int n, int m, const void* psi, const double* v, void* hpsi)
{
    return vloc_entry(c, lda, n, m, psi, v, hpsi, true, "qeb_vloc_psi_gamma");
}

int qeb_calbec(qeb_ctx* c, int n, const void* beta, int nkb, const void* psi, int lda, int m, void* becp)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (!c->have_k) QEB_FAIL(c, 70, "qeb_calbec: qeb_set_kpoint has not been called");
    if (n < 1 || n > c->npwx || lda < n || m < 0 || !psi || !becp) QEB_FAIL(c, 71, "qeb_calbec: bad arguments");
    if (!beta) {
        if (nkb != c->nkb) QEB_FAIL(c, 72, "qeb_calbec: nkb differs from the context's vkb");
        beta = c->vkb_d;
    }
    if (nkb == 0 || m == 0) return 0;
    int rc;
    const void *beta_d, *psi_d;
    if ((rc = stage_in(c, beta, (size_t)c->npwx * nkb * 16, c->beta_stage, &beta_d))) return rc;
    if ((rc = stage_in(c, psi, (size_t)lda * m * 16, c->psi_stage, &psi_d))) return rc;
    const size_t el = c->gamma ? 8 : 16;
    const bool keep = (beta_d == (const void*)c->vkb_d);
    void* out_d;
    if (keep) {
        if ((rc = ensure_becp(c, nkb, m))) return rc;
        out_d = c->becp_d.p;
    } else {
        QEB_CUDA(c, c->misc_stage.ensure((size_t)nkb * m * el));
        out_d = c->misc_stage.p;
    }
    if ((rc = run_calbec(c, (const cplx*)beta_d, c->npwx, nkb, (const cplx*)psi_d, lda, n, m, out_d))) return rc;
    if (keep) c->becp_m = m;
    QEB_CUDA(c, cudaMemcpyAsync(becp, out_d, (size_t)nkb * m * el,
                                is_device_ptr(becp) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, c->stream));
    if (!is_device_ptr(becp)) {
        QEB_CUDA(c, cudaStreamSynchronize(c->stream));
        return 0;
    }
    return finish(c);
}

int qeb_add_vuspsi(qeb_ctx* c, int lda, int n, int m, void* hpsi, const void* becp)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    int rc = check_call(c, lda, n, m, hpsi, hpsi, "qeb_add_vuspsi");
    if (rc) return rc;
    if (m == 0 || c->nkb == 0) return 0;
    const size_t el = c->gamma ? 8 : 16;
    if (becp) {
        if ((rc = ensure_becp(c, c->nkb, m))) return rc;
        QEB_CUDA(c, cudaMemcpyAsync(c->becp_d.p, becp, (size_t)c->nkb * m * el,
                                    is_device_ptr(becp) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
        c->becp_m = m;
    } else if (c->becp_m < m) {
        QEB_FAIL(c, 73, "qeb_add_vuspsi: no projections in the context for these bands");
    }
    const size_t bytes = (size_t)lda * m * 16;
    const bool out_dev = is_device_ptr(hpsi);
    const void* hin;
    if ((rc = stage_in(c, hpsi, bytes, c->hpsi_stage, &hin))) return rc;
    if ((rc = run_addvus(c, lda, n, m, (cplx*)hin))) return rc;
    if (!out_dev) {
        QEB_CUDA(c, cudaMemcpyAsync(hpsi, hin, bytes, cudaMemcpyDeviceToHost, c->stream));
        QEB_CUDA(c, cudaStreamSynchronize(c->stream));
        return 0;
    }
    return finish(c);
}

int qeb_set_qq(qeb_ctx* c, int nat, int nhm, const double* qq_at)
{
    if (!c) return 1;
    ++c->state_epoch;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (!c->have_nl) QEB_FAIL(c, 42, "qeb_set_qq: qeb_set_nonlocal has not been called");
    if (nat != c->nat || nhm != c->nhm) QEB_FAIL(c, 43, "qeb_set_qq: nat/nhm differ from qeb_set_nonlocal");
    const size_t nd = (size_t)nhm * nhm * nat;
    if (nd && !qq_at) QEB_FAIL(c, 43, "qeb_set_qq: null qq_at");
    QEB_CUDA(c, c->qq_d.ensure(nd * 8 + 8));
    if (nd)
        QEB_CUDA(c, cudaMemcpyAsync(c->qq_d.p, qq_at, nd * 8,
                                    is_device_ptr(qq_at) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
    QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    c->have_qq = true;
    return 0;
}

// s_psi_ (PW/src/s_psi.f90:79-262): spsi = psi (all lda rows, threaded_memcpy :118), then, when there are
// projectors with an augmentation part, spsi(1:n,:) += vkb * (blockdiag(qq_at) * becp) with the becp that h_psi or
// calbec left behind (:15, :132 "the becp are already computed"). Atoms of a non-ultrasoft type contribute
// ps = 0 (:232-240): the caller passes qq_at = 0 for them, which is what uspp holds for such types.
int qeb_s_psi(qeb_ctx* c, int lda, int n, int m, const void* psi, void* spsi)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    int rc = check_call(c, lda, n, m, psi, spsi, "qeb_s_psi");
    if (rc) return rc;
    if (m == 0) return 0;
    const size_t bytes = (size_t)lda * m * 16;
    const bool okvan = c->nkb > 0 && c->have_qq;
    if (okvan && c->becp_m < m) QEB_FAIL(c, 76, "qeb_s_psi: no projections in the context for these bands (h_psi/calbec first)");
    const bool out_dev = is_device_ptr(spsi);
    cplx* sd = (cplx*)spsi;
    if (!out_dev) {
        QEB_CUDA(c, c->hpsi_stage.ensure(bytes));
        sd = (cplx*)c->hpsi_stage.p;
    }
    QEB_CUDA(c, cudaMemcpyAsync(sd, psi, bytes, is_device_ptr(psi) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice,
                                c->stream));
    if (okvan) {
        const int ldps = (c->nkb + 1) & ~1;
        QEB_CUDA(c, cudaMemsetAsync(c->ps_d.p, 0, (size_t)ldps * m * (c->gamma ? 8 : 16), c->stream));
        if ((rc = run_addvus(c, lda, n, m, sd, 0, (const double*)c->qq_d.p))) return rc;
    }
    if (!out_dev) {
        QEB_CUDA(c, cudaMemcpyAsync(spsi, sd, bytes, cudaMemcpyDeviceToHost, c->stream));
        QEB_CUDA(c, cudaStreamSynchronize(c->stream));
        return 0;
    }
    return finish(c);
}

// s_psi_nc (PW/src/s_psi.f90:308-370): qq_nc(nhm,nhm,nat,4) complex, per atom: qq_so(:,:,1:4,ityp(na)) with spin-orbit,
// else qq_at(:,:,na) in blocks 1 and 4 and zero in 2 and 3 (:336-351)
int qeb_set_qq_nc(qeb_ctx* c, const void* qq_nc)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (!c->have_nl) QEB_FAIL(c, 51, "qeb_set_qq_nc: qeb_set_nonlocal has not been called (nat, nhm, nh, ofsbeta)");
    if (!qq_nc) QEB_FAIL(c, 40, "qeb_set_qq_nc: null qq_nc");
    const size_t nd = (size_t)c->nhm * c->nhm * c->nat * 4;
    QEB_CUDA(c, c->qnc_d.ensure(nd * 16 + 16));
    if (nd)
        QEB_CUDA(c, cudaMemcpyAsync(c->qnc_d.p, qq_nc, nd * 16,
                                    is_device_ptr(qq_nc) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
    QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    c->have_qnc = true;
    return 0;
}

// spsi(lda*2,m) = psi + vkb (qq_nc becp%nc) with the becp%nc the preceding qeb_h_psi_nc left in the context
int qeb_s_psi_nc(qeb_ctx* c, int lda, int n, int m, const void* psi, void* spsi)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    int rc = check_call(c, lda, n, m, psi, spsi, "qeb_s_psi_nc");
    if (rc) return rc;
    if (c->dist || c->gamma) QEB_FAIL(c, 63, "qeb_s_psi_nc: serial k-point descriptor only");
    if (m == 0) return 0;
    const size_t bytes = (size_t)lda * 2 * m * 16;
    const bool okvan = c->nkb > 0 && c->have_qnc;
    if (okvan && c->becp_m < 2 * m)
        QEB_FAIL(c, 76, "qeb_s_psi_nc: no projections in the context for these spinors (qeb_h_psi_nc first)");
    const bool out_dev = is_device_ptr(spsi);
    cplx* sd = (cplx*)spsi;
    if (!out_dev) {
        QEB_CUDA(c, c->hpsi_stage.ensure(bytes));
        sd = (cplx*)c->hpsi_stage.p;
    }
    QEB_CUDA(c, cudaMemcpyAsync(sd, psi, bytes, is_device_ptr(psi) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice,
                                c->stream));
    if (okvan) {
        const int nkb = c->nkb, ldps = (nkb + 1) & ~1;
        QEB_CUDA(c, cudaMemsetAsync(c->ps_d.p, 0, (size_t)ldps * 2 * m * 16, c->stream));
        QEB_CUDA(c, launch_apply_deeq_nc(c->becp_d.p, nkb, m, c->row_atom_d, c->ofs_d, c->nh_d, c->qnc_d.p, c->nhm, c->nat,
                                         c->ps_d.p, ldps, c->stream));
        QEB_CUDA(c, launch_addvus(c->vkb_d, c->npwx, c->ps_d.p, ldps, n, nkb, 2 * m, true, sd, lda, c->stream));
    }
    if (!out_dev) {
        QEB_CUDA(c, cudaMemcpyAsync(spsi, sd, bytes, cudaMemcpyDeviceToHost, c->stream));
        QEB_CUDA(c, cudaStreamSynchronize(c->stream));
        return 0;
    }
    return finish(c);
}

int qeb_get_becp(qeb_ctx* c, void* becp, int m)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (m > c->becp_m || !becp) QEB_FAIL(c, 74, "qeb_get_becp: no projections for that many bands");
    const size_t el = c->gamma ? 8 : 16;
    QEB_CUDA(c, cudaMemcpyAsync(becp, c->becp_d.p, (size_t)c->nkb * m * el,
                                is_device_ptr(becp) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, c->stream));
    QEB_CUDA(c, cudaStreamSynchronize(c->stream));
    return 0;
}

int qeb_nonlocal_only(qeb_ctx* c, int lda, int n, int m, const void* psi, void* hpsi)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    int rc = check_call(c, lda, n, m, psi, hpsi, "qeb_nonlocal_only");
    if (rc) return rc;
    if (!is_device_ptr(psi) || !is_device_ptr(hpsi)) QEB_FAIL(c, 75, "qeb_nonlocal_only: device pointers required");
    if ((rc = h_psi_cols(c, lda, n, m, (const cplx*)psi, (cplx*)hpsi, false, true, false))) return rc;
    return finish(c);
}

int qeb_fft(qeb_ctx* c, int sign, int kind, void* f, int howmany)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (!c->have_grid) QEB_FAIL(c, 80, "qeb_fft: qeb_set_grid has not been called");
    if (!f || howmany < 1 || sign == 0 || (kind != QEB_FFT_RHO && kind != QEB_FFT_WAVE)) QEB_FAIL(c, 81, "qeb_fft: bad arguments");
    const size_t nnr = (size_t)c->nr1 * c->nr2 * c->nr3;
    const size_t bytes = nnr * 16 * howmany;
    const bool dev = is_device_ptr(f);
    const void* fin;
    int rc;
    if ((rc = stage_in(c, f, bytes, c->hpsi_stage, &fin))) return rc;
    QEB_CUDA(c, fullfft_launch(c->fpx, c->fpy, c->fpz, (cplx*)fin, howmany, sign, c->max_smem, c->stream));
    if (!dev) {
        QEB_CUDA(c, cudaMemcpyAsync(f, fin, bytes, cudaMemcpyDeviceToHost, c->stream));
        QEB_CUDA(c, cudaStreamSynchronize(c->stream));
        return 0;
    }
    return finish(c);
}

int qeb_divide_all(int m, int nproc, int me, int* start, int* end)
{
    if (m < 0 || nproc < 1 || me < 0 || me >= nproc || !start || !end) return 1;
    const int base = m / nproc, rest = m % nproc;
    // first `rest` groups get base+1 columns (UtilXlib/divide.f90:55-106)
    const int s = me * base + (me < rest ? me : rest);
    *start = s;
    *end = s + base + (me < rest ? 1 : 0);
    return 0;
}

int qeb_set_nccl(qeb_ctx* c, void* comm, int nranks, int rank)
{
    return qeb_set_nccl_comm(c, QEB_COMM_BGRP, comm, nranks, rank);
}

int qeb_set_nccl_comm(qeb_ctx* c, int role, void* comm, int nranks, int rank)
{
    if (!c || nranks < 1 || rank < 0 || rank >= nranks || (role != QEB_COMM_BGRP && role != QEB_COMM_RG)) return 1;
    ++c->state_epoch;
    const NcclApi& A = nccl_api();
    if (role == QEB_COMM_BGRP) {
        if (c->nccl && c->nccl_owned && A.ok) A.CommDestroy(c->nccl);
        c->nccl = comm;
        c->nccl_owned = false;
        c->nranks = nranks;
        c->rank = rank;
    } else {
        if (c->nccl_rg && c->nccl_rg_owned && A.ok) A.CommDestroy(c->nccl_rg);
        c->nccl_rg = comm;
        c->nccl_rg_owned = false;
    }
    return 0;
}

int qeb_nccl_unique_id(void* id128)
{
    const NcclApi& A = nccl_api();
    if (!A.ok || !id128) return 1;
    return A.GetUniqueId((NcclUniqueId*)id128);
}

int qeb_nccl_init(qeb_ctx* c, int role, const void* id128, int nranks, int rank)
{
    if (!c) return 1;
    if (!id128 || nranks < 1 || rank < 0 || rank >= nranks || (role != QEB_COMM_BGRP && role != QEB_COMM_RG))
        QEB_FAIL(c, 91, "qeb_nccl_init: bad arguments");
    QEB_CUDA(c, cudaSetDevice(c->device));
    const NcclApi& A = nccl_api();
    if (!A.ok) QEB_FAIL(c, 92, std::string("qeb_nccl_init: ") + A.why);
    NcclUniqueId id;
    memcpy(&id, id128, sizeof(id));
    NcclComm comm = nullptr;
    const int rc = A.CommInitRank(&comm, nranks, id, rank);
    if (rc) QEB_FAIL(c, 900 + rc, std::string("ncclCommInitRank: ") + A.GetErrorString(rc));
    qeb_set_nccl_comm(c, role, comm, nranks, rank);
    if (role == QEB_COMM_BGRP)
        c->nccl_owned = true;
    else
        c->nccl_rg_owned = true;
    return 0;
}

int qeb_h_psi_bgrp(qeb_ctx* c, int lda, int n, int m, const void* psi, void* hpsi, int nbgrp, int my_bgrp)
{
    if (!c) return 1;
    int s, e;
    if (qeb_divide_all(m, nbgrp, my_bgrp, &s, &e)) QEB_FAIL(c, 90, "qeb_h_psi_bgrp: bad band-group arguments");
    const bool gather = c->nccl && c->nranks == nbgrp && c->rank == my_bgrp && nbgrp > 1 && is_device_ptr(hpsi);
    if (e > s) {
        const char* p = (const char*)psi + (size_t)s * lda * 16;
        char* h = (char*)hpsi + (size_t)s * lda * 16;
        const bool was_async = c->async;
        if (gather) c->async = true;  // the all-gather is queued behind the kernels; one sync at the end
        const int rc = qeb_h_psi(c, lda, n, e - s, p, h);
        c->async = was_async;
        if (rc) return rc;
    }
    if (!gather) return 0;
    // mp_allgather(hpsi, column_type, recv_counts, displs, inter_bgrp_comm)  (PW/src/h_psi.f90:58-63): every group
    // sends its column block to all others; blocks may differ by one column (divide_all)
    const NcclApi& A = nccl_api();
    if (!A.ok) QEB_FAIL(c, 92, std::string("qeb_h_psi_bgrp: ") + A.why);
    int rc = A.GroupStart();
    for (int q = 0; q < nbgrp && rc == 0; ++q) {
        if (q == my_bgrp) continue;
        int qs, qe;
        qeb_divide_all(m, nbgrp, q, &qs, &qe);
        if (e > s) rc = A.Send((char*)hpsi + (size_t)s * lda * 16, (size_t)(e - s) * lda * 2, kNcclDouble, q, c->nccl, c->stream);
        if (rc == 0 && qe > qs)
            rc = A.Recv((char*)hpsi + (size_t)qs * lda * 16, (size_t)(qe - qs) * lda * 2, kNcclDouble, q, c->nccl, c->stream);
    }
    const int rc2 = A.GroupEnd();
    if (rc == 0) rc = rc2;
    if (rc) QEB_FAIL(c, 900 + rc, std::string("all-gather of hpsi: ") + A.GetErrorString(rc));
    return finish(c);
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
