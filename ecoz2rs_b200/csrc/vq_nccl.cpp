// vq_nccl.cpp -- run-time binding of NCCL (see vq_nccl.h)
#include "vq_nccl.h"

#include <dlfcn.h>

#include <cstdlib>

#include <mutex>
#include <string>

#include "vq_engine.h"

namespace vqb200 {

static NcclApi g_api;
static bool g_loaded = false, g_failed = false;
static std::mutex g_mu;

const NcclApi *nccl_api() {
    std::lock_guard<std::mutex> lock(g_mu);
    if (g_loaded) return &g_api;
    if (g_failed) return nullptr;
    // RTLD_NOLOAD first: reuse the NCCL torch already loaded in this process.  The loader keeps ONE
    // library per soname: if a process imports torch AFTER this point, torch gets the NCCL loaded
    // here, so ECOZ2_VQ_NCCL_LIB can name the one to load (the ctypes mirror pre-loads torch's
    // bundled copy, ecoz2rs_b200/vq.py).
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!h) {
        const char *path = std::getenv("ECOZ2_VQ_NCCL_LIB");
        if (path && *path) h = dlopen(path, RTLD_NOW | RTLD_GLOBAL);
    }
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) {
        g_failed = true;
        set_last_error(std::string("cannot load libnccl: ") + dlerror());
        return nullptr;
    }
    bool ok = true;
    auto sym = [&](const char *name) {
        void *p = dlsym(h, name);
        if (!p) ok = false;
        return p;
    };
    g_api.GetUniqueId = (int (*)(NcclUniqueId *)) sym("ncclGetUniqueId");
    g_api.CommInitRank = (int (*)(void **, int, NcclUniqueId, int)) sym("ncclCommInitRank");
    g_api.CommInitAll = (int (*)(void **, int, const int *)) sym("ncclCommInitAll");
    g_api.CommDestroy = (int (*)(void *)) sym("ncclCommDestroy");
    g_api.AllReduce = (int (*)(const void *, void *, size_t, int, int, void *, cudaStream_t)) sym(
        "ncclAllReduce");
    g_api.GroupStart = (int (*)()) sym("ncclGroupStart");
    g_api.GroupEnd = (int (*)()) sym("ncclGroupEnd");
    g_api.GetErrorString = (const char *(*) (int) ) sym("ncclGetErrorString");
    if (!ok) {
        g_failed = true;
        set_last_error("libnccl is missing a required symbol");
        return nullptr;
    }
    g_loaded = true;
    return &g_api;
}

}  // namespace vqb200
