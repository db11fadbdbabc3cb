// nccl_dyn.h -- NCCL bound at run time (dlopen), not at link time.
//
// libmrisim.so is loaded into processes that already carry an NCCL (PyTorch bundles its own
// libnccl.so.2) and into plain C++ programs that do not (the drop-in gpu_sim.out).  Linking a second
// copy would put two NCCLs with the same SONAME into one process, so the library is resolved when
// the first communicator is asked for: an NCCL already mapped into the process wins (RTLD_NOLOAD),
// then $MRISIM_NCCL_LIB, then the system's libnccl.so.2.  Only the eight entry points the path
// needs are bound; <nccl.h> supplies the types.
#pragma once
#include <dlfcn.h>
#include <nccl.h>

#include <cstdlib>
#include <mutex>
#include <string>

namespace mrisim_nccl {

struct Api {
    ncclResult_t (*GetVersion)(int*) = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    void* handle = nullptr;
    std::string error, origin;
};

inline Api& api() {
    static Api a;
    static std::once_flag once;
    std::call_once(once, [] {
        const char* env = std::getenv("MRISIM_NCCL_LIB");
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
        if (h) a.origin = "already loaded in the process";
        if (!h && env && *env) { h = dlopen(env, RTLD_NOW | RTLD_GLOBAL); if (h) a.origin = env; }
        if (!h) { h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL); if (h) a.origin = "libnccl.so.2 (system search path)"; }
        if (!h) { a.error = std::string("NCCL not found: ") + (dlerror() ? dlerror() : "dlopen failed") + " (set MRISIM_NCCL_LIB)"; return; }
        a.handle = h;
        auto bind = [&](const char* name) { void* f = dlsym(h, name); if (!f && a.error.empty()) a.error = std::string("NCCL symbol missing: ") + name; return f; };
        a.GetVersion = reinterpret_cast<decltype(a.GetVersion)>(bind("ncclGetVersion"));
        a.GetUniqueId = reinterpret_cast<decltype(a.GetUniqueId)>(bind("ncclGetUniqueId"));
        a.CommInitRank = reinterpret_cast<decltype(a.CommInitRank)>(bind("ncclCommInitRank"));
        a.CommInitAll = reinterpret_cast<decltype(a.CommInitAll)>(bind("ncclCommInitAll"));
        a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(bind("ncclCommDestroy"));
        a.AllReduce = reinterpret_cast<decltype(a.AllReduce)>(bind("ncclAllReduce"));
        a.GroupStart = reinterpret_cast<decltype(a.GroupStart)>(bind("ncclGroupStart"));
        a.GroupEnd = reinterpret_cast<decltype(a.GroupEnd)>(bind("ncclGroupEnd"));
        a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(bind("ncclGetErrorString"));
    });
    return a;
}

inline bool ok() { return api().handle && api().error.empty(); }

}  // namespace mrisim_nccl
