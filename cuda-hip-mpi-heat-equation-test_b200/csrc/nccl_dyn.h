// nccl_dyn.h — run-time binding to NCCL (dlopen), so libheat.so loads on machines without NCCL
// and uses whichever libnccl.so.2 the host process has already mapped (e.g. torch's bundled one).
// Only the handful of entry points the halo exchange needs; types restated from the stable NCCL ABI.
#pragma once
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <stddef.h>

namespace heat {

typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;                       // 0 == ncclSuccess
enum { ncclInt8 = 0, ncclFloat32 = 7, ncclFloat64 = 8 };

struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int*) = nullptr;
    bool ok = false;
};

inline const NcclApi& nccl_api() {
    static NcclApi api = [] {
        NcclApi a;
        const char* names[] = {getenv("HEAT_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
        for (const char* n : names) {
            if (!n || !*n) continue;
            a.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
            if (a.lib) break;
        }
        if (!a.lib) return a;
#define HEAT_NCCL_SYM(field, name) \
        *(void**)(&a.field) = dlsym(a.lib, name); if (!a.field) return a;
        HEAT_NCCL_SYM(GetUniqueId, "ncclGetUniqueId")
        HEAT_NCCL_SYM(CommInitRank, "ncclCommInitRank")
        HEAT_NCCL_SYM(CommDestroy, "ncclCommDestroy")
        HEAT_NCCL_SYM(Send, "ncclSend")
        HEAT_NCCL_SYM(Recv, "ncclRecv")
        HEAT_NCCL_SYM(GroupStart, "ncclGroupStart")
        HEAT_NCCL_SYM(GroupEnd, "ncclGroupEnd")
        HEAT_NCCL_SYM(GetErrorString, "ncclGetErrorString")
        HEAT_NCCL_SYM(GetVersion, "ncclGetVersion")
#undef HEAT_NCCL_SYM
        a.ok = true;
        return a;
    }();
    return api;
}

}  // namespace heat
