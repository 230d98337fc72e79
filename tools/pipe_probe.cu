// Micro-probe: do FP64 vector instructions (DFMA) and FP64 tensor instructions (DMMA.8x8x4) share an execution pipe on
// sm_100a? Times (a) a CTA of DFMA-only warps, (b) DMMA-only warps, (c) both kinds of warps resident together.
// If (c) ~ max(a, b) the pipes are independent and the V_loc FFT passes can run under the projector GEMMs.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/pipe_probe.bin tools/pipe_probe.cu
#include <cuda_runtime.h>
#include <stdio.h>

__device__ __forceinline__ void dmma(double& d0, double& d1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(d0), "+d"(d1)
                 : "d"(a), "d"(b));
}

// mode bit 0: warps [0, nv) run DFMA chains; bit 1: warps [nv, nv+nt) run DMMA chains
__global__ void __launch_bounds__(512) k_probe(int iters, int nv, int nt, int mode, double* out)
{
    const int w = threadIdx.x / 32;
    double acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = threadIdx.x * 1e-3 + i;
    const double a = 1.0000001, b = 1e-9;
    if (w < nv) {
        if (!(mode & 1)) return;
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 8; ++i) acc[i] = fma(acc[i], a, b);
        }
    } else if (w < nv + nt) {
        if (!(mode & 2)) return;
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 4; ++i) dmma(acc[2 * i], acc[2 * i + 1], a, b);
        }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += acc[i];
    if (s == 123.456) out[0] = s;
}

static float run(int iters, int nv, int nt, int mode, double* out)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int grid = 148 * 4;
    k_probe<<<grid, (nv + nt) * 32>>>(iters, nv, nt, mode, out);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k_probe<<<grid, (nv + nt) * 32>>>(iters, nv, nt, mode, out);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms;
}

int main()
{
    double* out;
    cudaMalloc(&out, 64);
    const int iters = 20000;
    for (int nv = 4; nv <= 8; nv += 4)
        for (int nt = 4; nt <= 8; nt += 4) {
            const float a = run(iters, nv, nt, 1, out), b = run(iters, nv, nt, 2, out), c = run(iters, nv, nt, 3, out);
            // per SM: 4 CTAs resident x nv warps x iters x 8 DFMA; DMMA: 4 per iteration, 256 FMA each
            const double dfma_tf = 2.0 * 148 * 4 * nv * 32.0 * iters * 8 / (a * 1e-3) / 1e12;
            const double dmma_tf = 2.0 * 148 * 4 * nt * 256.0 * iters * 4 / (b * 1e-3) / 1e12;
            printf("{\"probe\":\"fp64_pipes\",\"vector_warps\":%d,\"tensor_warps\":%d,\"ms_dfma\":%.3f,\"ms_dmma\":%.3f,"
                   "\"ms_both\":%.3f,\"dfma_tflops\":%.2f,\"dmma_tflops\":%.2f,\"both_over_max\":%.3f,\"both_over_sum\":%.3f}\n",
                   nv, nt, a, b, c, dfma_tf, dmma_tf, c / (a > b ? a : b), c / (a + b));
        }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        printf("error %s\n", cudaGetErrorString(e));
        return 1;
    }
    return 0;
}
