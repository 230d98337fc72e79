// libqeb200, translation unit 3 of 6: h_psi on HOST arrays - column chunks through three streams (copy-in, kernels,
// copy-out), with the first and last chunks sub-chunked.
#include "context.h"

extern "C" {

// h_psi with HOST psi and hpsi: column chunks flow through three streams, so the H2D copy of chunk i+1 and the D2H
// copy of chunk i-1 overlap the kernels of chunk i. Every column is independent in h_psi (h_psi.f90:53-67 splits
// columns the same way over band groups), so the result is the one of the single call.
int h_psi_host_pipeline(qeb_ctx* c, int lda, int n, int m, const void* psi, void* hpsi)
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

}  // extern "C"
