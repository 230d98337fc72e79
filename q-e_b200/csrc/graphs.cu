// libqeb200, translation unit 4 of 6: the launch-bound regime - CUDA-graph replay of repeated h_psi calls and
// qeb_h_psi_multi (the k-points of a pool captured into one graph).
#include "context.h"

extern "C" {

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

bool graph_wanted(qeb_ctx* c, int n, int m)
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
int h_psi_graph(qeb_ctx* c, int lda, int n, int m, const cplx* psi_d, cplx* hpsi_d, bool* done)
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

}  // extern "C"
