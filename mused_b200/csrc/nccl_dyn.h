// NCCL entry points resolved at run time (dlopen of the libnccl the host process already uses --
// torch's bundled one): the library must load on a machine without NCCL (CPU build check, symbol
// test) and must not pin a second NCCL copy next to torch's.  Only the column-sharded mode calls it.
#pragma once
#include <dlfcn.h>
#include <nccl.h>

namespace mused {

struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool ok() const { return handle && GetUniqueId && CommInitRank && CommDestroy && AllReduce && GetErrorString; }
};

inline NcclApi& nccl_api() {
    static NcclApi api;
    return api;
}

// path = NULL: whatever "libnccl.so.2" the process resolves (already mapped when torch initialised NCCL)
inline bool nccl_load(const char* path, std::string* err) {
    NcclApi& a = nccl_api();
    if (a.ok()) return true;
    const char* names[] = {path, "libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
        if (!n || !*n) continue;
        a.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (a.handle) break;
    }
    if (!a.handle) {
        if (err) *err = std::string("dlopen(libnccl) failed: ") + (dlerror() ? dlerror() : "?");
        return false;
    }
    a.GetUniqueId = (decltype(a.GetUniqueId))dlsym(a.handle, "ncclGetUniqueId");
    a.CommInitRank = (decltype(a.CommInitRank))dlsym(a.handle, "ncclCommInitRank");
    a.CommDestroy = (decltype(a.CommDestroy))dlsym(a.handle, "ncclCommDestroy");
    a.AllReduce = (decltype(a.AllReduce))dlsym(a.handle, "ncclAllReduce");
    a.GetErrorString = (decltype(a.GetErrorString))dlsym(a.handle, "ncclGetErrorString");
    if (!a.ok()) {
        if (err) *err = "libnccl lacks an expected symbol";
        return false;
    }
    return true;
}

#define MUSED_NCCL(call)                                                                             \
    do {                                                                                             \
        ncclResult_t _r = (call);                                                                    \
        if (_r != ncclSuccess) {                                                                     \
            mused::set_error("%s failed: %s (%s:%d)", #call, nccl_api().GetErrorString(_r), __FILE__, __LINE__); \
            return MUSED_ERR_CUDA;                                                                   \
        }                                                                                            \
    } while (0)

}  // namespace mused
