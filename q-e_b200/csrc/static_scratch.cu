// qeb_calbec_static: calbec (Modules/becmod.f90:138-172) for callers that cannot allocate - code under CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc. It is the projector kernel of qeb_calbec (gemm.cu: k_calbec3m /
// k_calbec + k_calbec_reduce) with its split-K partial-sum workspace taken from a file-scope device array instead of the
// context's grow-on-demand DevBuf. No counterpart in the reference.
//
// The array occupies QEB_STATIC_SCRATCH_BYTES of HBM on every GPU that loads the library, until the process exits. All
// calls on a device share it and it keeps the previous call's contents: the part a call uses is cleared first, and two
// calls that overlap on the GPU (different streams / contexts of one device) race on it - serialise them.
//
// In host code `g_static_scratch` is only the host-side stand-in of the symbol, not a device address: the address comes
// from cudaGetSymbolAddress, once per device with that device current (static_scratch_ptr below). Device code in this
// file (k_static_scratch_clear) names the array directly.
#include <mutex>
#include "context.h"
#include "gemm.h"

#define QEB_STATIC_SCRATCH_BYTES QEB_STATIC_SCRATCH_MAX_BYTES /* include/qeb200.h: 32 MiB */

static __device__ __align__(256) unsigned char g_static_scratch[QEB_STATIC_SCRATCH_BYTES];

namespace {

constexpr int kMaxDevices = 64;
void* g_scratch_addr[kMaxDevices] = {};
std::mutex g_scratch_mu;

// zero the first `bytes` (a multiple of 16) of the shared array
__global__ void k_static_scratch_clear(size_t bytes)
{
    uint4* p = reinterpret_cast<uint4*>(g_static_scratch);
    const size_t n = bytes / 16;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        p[i] = make_uint4(0u, 0u, 0u, 0u);
}

// device address of the array on `device`, which must be the current device
cudaError_t static_scratch_ptr(int device, void** out)
{
    if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lk(g_scratch_mu);
    if (!g_scratch_addr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess) return e;
        g_scratch_addr[device] = p;
    }
    *out = g_scratch_addr[device];
    return cudaSuccess;
}

}  // namespace

extern "C" size_t qeb_static_scratch_bytes(void) { return QEB_STATIC_SCRATCH_BYTES; }

extern "C" int qeb_calbec_static(qeb_ctx* c, int n, const void* beta, int nkb, const void* psi, int lda, int m, void* becp)
{
    if (!c) return 1;
    QEB_CUDA(c, cudaSetDevice(c->device));
    if (!c->have_k) QEB_FAIL(c, 70, "qeb_calbec_static: qeb_set_kpoint has not been called");
    if (n < 1 || n > c->npwx || lda < n || m < 0 || nkb < 0 || !psi || !becp) QEB_FAIL(c, 71, "qeb_calbec_static: bad arguments");
    if (c->dist && c->DH.nproc > 1)
        QEB_FAIL(c, 74, "qeb_calbec_static: not available on a slab-distributed context (the R&G sum of becp is qeb_calbec's)");
    if (!beta) {
        if (nkb != c->nkb) QEB_FAIL(c, 72, "qeb_calbec_static: nkb differs from the context's vkb");
        beta = c->vkb_d;
    }
    if (nkb == 0 || m == 0) return 0;
    // nothing is staged here (staging buffers are allocations): every array is the caller's and lives in HBM
    if (!is_device_ptr(beta) || !is_device_ptr(psi) || !is_device_ptr(becp))
        QEB_FAIL(c, 93, "qeb_calbec_static: beta, psi and becp must be device pointers");
    const bool cplx_mode = !c->gamma;
    const size_t one = calbec_workspace_bytes(nkb, m, 1, cplx_mode);
    if (one > QEB_STATIC_SCRATCH_BYTES)
        QEB_FAIL(c, 75, "qeb_calbec_static: nkb*m = " + std::to_string((size_t)nkb * m) + " projections need " +
                            std::to_string(one) + " bytes of scratch, beyond the static maximum of " +
                            std::to_string((size_t)QEB_STATIC_SCRATCH_BYTES) + " bytes (use qeb_calbec or fewer bands per call)");
    int nsplit, gchunk;
    calbec_plan(n, nkb, m, cplx_mode, c->num_sms, &nsplit, &gchunk);
    const int fit = (int)(QEB_STATIC_SCRATCH_BYTES / one);
    if (nsplit > fit) {  // fewer, longer splits so that the partial sums fit the array
        gchunk = ((n + fit - 1) / fit + 15) / 16 * 16;
        nsplit = (n + gchunk - 1) / gchunk;
    }
    const size_t bytes = calbec_workspace_bytes(nkb, m, nsplit, cplx_mode);
    void* scratch = nullptr;
    QEB_CUDA(c, static_scratch_ptr(c->device, &scratch));
    int blocks = (int)((bytes / 16 + 255) / 256);
    if (blocks > c->num_sms * 8) blocks = c->num_sms * 8;
    if (blocks < 1) blocks = 1;
    k_static_scratch_clear<<<blocks, 256, 0, c->stream>>>((bytes + 15) / 16 * 16);
    QEB_CUDA(c, cudaGetLastError());
    QEB_CUDA(c, launch_calbec((const cplx*)beta, c->npwx, (const cplx*)psi, lda, n, nkb, m, cplx_mode, c->gstart, nsplit, gchunk,
                              scratch, becp, c->stream));
    return finish(c);
}
