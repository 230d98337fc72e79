// libqeb200, translation unit 5 of 6: one process per GPU - NCCL communicators by role, the peer-to-peer mapping of the
// R&G exchange buffers (cudaIpc), divide_all and h_psi over band groups.
#include "context.h"

extern "C" {

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

}  // extern "C"
