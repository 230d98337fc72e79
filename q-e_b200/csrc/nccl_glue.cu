#include "nccl_glue.h"
#include <dlfcn.h>
#include <stdlib.h>

namespace qeb {

static NcclApi g_api;
static bool g_tried = false;

const NcclApi& nccl_api()
{
    if (g_tried) return g_api;
    g_tried = true;
    void* h = nullptr;
    if (const char* p = getenv("QEB_NCCL_LIB")) h = dlopen(p, RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) {
        g_api.why = "libnccl.so.2 not found (set QEB_NCCL_LIB)";
        return g_api;
    }
#define QEB_SYM(field, name)                                              \
    *(void**)(&g_api.field) = dlsym(h, name);                             \
    if (!g_api.field) {                                                   \
        g_api.why = "symbol " name " missing in libnccl";                 \
        return g_api;                                                     \
    }
    QEB_SYM(GetUniqueId, "ncclGetUniqueId")
    QEB_SYM(CommInitRank, "ncclCommInitRank")
    QEB_SYM(CommDestroy, "ncclCommDestroy")
    QEB_SYM(GroupStart, "ncclGroupStart")
    QEB_SYM(GroupEnd, "ncclGroupEnd")
    QEB_SYM(Send, "ncclSend")
    QEB_SYM(Recv, "ncclRecv")
    QEB_SYM(AllReduce, "ncclAllReduce")
    QEB_SYM(AllGather, "ncclAllGather")
    QEB_SYM(GetErrorString, "ncclGetErrorString")
#undef QEB_SYM
    g_api.ok = true;
    return g_api;
}

}  // namespace qeb
