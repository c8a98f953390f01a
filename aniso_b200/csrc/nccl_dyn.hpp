// NCCL is bound at run time (dlopen) so that the library also loads on a CPU-only box for the host-side
// entry points, and so that it shares the libnccl.so.2 already mapped by the host process (e.g. torch's).
// Only used when a handle is created with nranks > 1.
#pragma once
#include <dlfcn.h>
#include <nccl.h>

namespace an {

struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId*);
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
    ncclResult_t (*CommDestroy)(ncclComm_t);
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t);
    ncclResult_t (*GroupStart)();
    ncclResult_t (*GroupEnd)();
    const char* (*GetErrorString)(ncclResult_t);
};

inline NcclApi* load_nccl() {
    static NcclApi api;
    static bool tried = false, ok = false;
    if (tried) return ok ? &api : nullptr;
    tried = true;
    void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) return nullptr;
#define ANISO_NCCL_SYM(field, name)                                   \
    api.field = reinterpret_cast<decltype(api.field)>(dlsym(lib, name)); \
    if (!api.field) return nullptr;
    ANISO_NCCL_SYM(GetUniqueId, "ncclGetUniqueId")
    ANISO_NCCL_SYM(CommInitRank, "ncclCommInitRank")
    ANISO_NCCL_SYM(CommDestroy, "ncclCommDestroy")
    ANISO_NCCL_SYM(AllReduce, "ncclAllReduce")
    ANISO_NCCL_SYM(Broadcast, "ncclBroadcast")
    ANISO_NCCL_SYM(AllGather, "ncclAllGather")
    ANISO_NCCL_SYM(GroupStart, "ncclGroupStart")
    ANISO_NCCL_SYM(GroupEnd, "ncclGroupEnd")
    ANISO_NCCL_SYM(GetErrorString, "ncclGetErrorString")
#undef ANISO_NCCL_SYM
    ok = true;
    return &api;
}

}  // namespace an
