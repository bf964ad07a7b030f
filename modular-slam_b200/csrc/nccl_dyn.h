// nccl_dyn.h -- NCCL bound at run time (dlopen "libnccl.so.2"), so that libmslam_b200.so itself has no NCCL dependency:
// the single-GPU entry points load and run where NCCL is absent, and in a process that already holds an NCCL (PyTorch's
// bundled one) the SAME library instance is picked up by soname.  Only the sharded keyframe-database entry points
// (mslam_b200_db_attach* / _db_query_sharded_dev) touch this.
#pragma once
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <stddef.h>
#include <stdio.h>
#include <stdlib.h>

namespace mslam_b200
{
// the slice of nccl.h that is used (types as NCCL 2.x declares them; the ABI of these entry points is stable across 2.x)
typedef struct ncclComm* ncclComm_t;
typedef struct
{
    char internal[128];
} ncclUniqueId;
typedef int ncclResult_t;     // 0 = ncclSuccess
typedef int ncclDataType_t;   // ncclInt32 = 2
constexpr int kNcclInt32 = 2;

struct NcclApi
{
    void* handle = nullptr;
    ncclResult_t (*GetVersion)(int*) = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*CommAbort)(ncclComm_t) = nullptr;
    ncclResult_t (*CommCount)(const ncclComm_t, int*) = nullptr;
    ncclResult_t (*CommUserRank)(const ncclComm_t, int*) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    char why[256] = {0};

    bool ok() const { return handle != nullptr; }
};

// process-wide, loaded on first use (MSLAM_B200_NCCL_LIB overrides the library name)
inline NcclApi& nccl_api()
{
    static NcclApi api;
    static bool tried = false;
    if(tried)
        return api;
    tried = true;
    const char* names[] = {getenv("MSLAM_B200_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    void* h = nullptr;
    for(const char* n : names)
        if(n && (h = dlopen(n, RTLD_NOW | RTLD_GLOBAL)))
            break;
    if(!h)
    {
        snprintf(api.why, sizeof(api.why), "NCCL not found: %s", dlerror());
        return api;
    }
#define MSLAM_NCCL_SYM(field, name)                                                                                    \
    if(!(api.field = reinterpret_cast<decltype(api.field)>(dlsym(h, name))))                                           \
    {                                                                                                                  \
        snprintf(api.why, sizeof(api.why), "NCCL symbol %s missing", name);                                            \
        return api;                                                                                                    \
    }
    MSLAM_NCCL_SYM(GetVersion, "ncclGetVersion")
    MSLAM_NCCL_SYM(GetUniqueId, "ncclGetUniqueId")
    MSLAM_NCCL_SYM(CommInitRank, "ncclCommInitRank")
    MSLAM_NCCL_SYM(CommDestroy, "ncclCommDestroy")
    MSLAM_NCCL_SYM(CommAbort, "ncclCommAbort")
    MSLAM_NCCL_SYM(CommCount, "ncclCommCount")
    MSLAM_NCCL_SYM(CommUserRank, "ncclCommUserRank")
    MSLAM_NCCL_SYM(AllGather, "ncclAllGather")
    MSLAM_NCCL_SYM(GroupStart, "ncclGroupStart")
    MSLAM_NCCL_SYM(GroupEnd, "ncclGroupEnd")
    MSLAM_NCCL_SYM(GetErrorString, "ncclGetErrorString")
#undef MSLAM_NCCL_SYM
    api.handle = h;
    return api;
}
} // namespace mslam_b200
