// libqeb200, translation unit 2 of 6: the reference entry points - h_psi (PW/src/h_psi.f90:9-80: kinetic :138-147, V_loc
// :180/:220, nonlocal :229-236, Gamma fix :294-295), vloc_psi_k/gamma/nc, calbec, add_vuspsi, s_psi, hs_psi, the spinor
// variants, invfft/fwfft, the sum_band density loop - and the orchestration of the V_loc passes and projector GEMMs
// behind them (run_vloc incl. the slab-distributed sequence with its exchanges, run_calbec, run_addvus).
#include "context.h"

extern "C" {

// ---- internal pieces ----------------------------------------------------
int run_vloc(qeb_ctx* c, int lda, int n, int m, const cplx* psi, cplx* hpsi, bool kinetic)
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

int ensure_becp(qeb_ctx* c, int nkb, int m)
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

int run_calbec(qeb_ctx* c, const cplx* beta, int npwx, int nkb, const cplx* psi, int lda, int n, int m, void* becp_out)
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

int run_addvus(qeb_ctx* c, int lda, int n, int m, cplx* hpsi, int col0, const double* dmat)
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

int check_call(qeb_ctx* c, int lda, int n, int m, const void* psi, const void* hpsi, const char* who)
{
    if (!c->have_k) QEB_FAIL(c, 60, std::string(who) + ": qeb_set_kpoint has not been called");
    if (n != c->npw) QEB_FAIL(c, 61, std::string(who) + ": n differs from the npw of qeb_set_kpoint");
    if (lda < n || m < 0 || !psi || !hpsi) QEB_FAIL(c, 62, std::string(who) + ": bad arguments");
    return 0;
}

// columns [col0, col0+m) of a call over m_total columns; psi_d / hpsi_d point at column col0. The projections of
// the whole call live in becp_d (column j at j*nkb): the caller runs ensure_becp(m_total) once when col0 >= 0 is
// used with m_total > m.
int h_psi_cols(qeb_ctx* c, int lda, int n, int m, const cplx* psi_d, cplx* hpsi_d, bool do_vloc, bool do_nl,
                      bool kinetic, int col0, int m_total)
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


// sum_band, density part (PW/src/sum_band.f90:326-437 Gamma, :440-647 k): rho(r) += sum_b w_b |psi_b(r)|^2 for the
// m bands of evc, w_b = wg(b,ik)/omega. The reference loops wave_g2r + get_rho_k/get_rho_gamma per band (group);
// here: scatter + z pass, y pass, then one kernel that x-transforms every box of a batch and accumulates.
int check_call(qeb_ctx* c, int lda, int n, int m, const void* psi, const void* hpsi, const char* who);
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

}  // extern "C"
