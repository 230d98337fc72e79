// Runtime binding to NCCL (dlopen): the library has no link-time dependency on libnccl, so single-GPU users need
// no NCCL at all, and a process that already has a libnccl.so.2 loaded (e.g. PyTorch's bundled one) shares it.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

namespace qeb {

struct NcclUniqueId { char internal[128]; };  // ncclUniqueId (nccl.h: NCCL_UNIQUE_ID_BYTES = 128)
typedef void* NcclComm;                       // ncclComm_t

struct NcclApi {
    bool ok = false;
    const char* why = "";
    int (*GetUniqueId)(NcclUniqueId*) = nullptr;
    int (*CommInitRank)(NcclComm*, int, NcclUniqueId, int) = nullptr;
    int (*CommDestroy)(NcclComm) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    int (*Send)(const void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, NcclComm, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
constexpr int kNcclDouble = 8;  // ncclFloat64
constexpr int kNcclSum = 0;     // ncclSum

// loads libnccl on first use: $QEB_NCCL_LIB, else an already loaded libnccl.so.2, else the system one
const NcclApi& nccl_api();

}  // namespace qeb
