#pragma once
#include <cuda_runtime.h>
#include "common.cuh"
namespace qeb {
cudaError_t fullfft_launch(const FftPlan& px, const FftPlan& py, const FftPlan& pz, cplx* f, int howmany, int sign,
                           size_t max_smem, cudaStream_t st);
}
