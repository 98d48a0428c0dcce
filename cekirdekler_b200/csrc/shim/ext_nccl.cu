// ext_nccl.cu — NCCL broadcast over NVLink/NVSwitch for whole-array `read` inputs.
//
// The reference uploads a whole-`read` array from the host to EVERY device (Worker.cs:833-834, called per worker from
// Cores.cs:747-768).  Here the array crosses the host link once, into one device (the root rotates from call to call so
// that no single host link carries every upload), and NCCL broadcasts it from there to the other devices of the
// cruncher over NVLink.  One process drives all devices (the reference's execution model), so the communicators come
// from ncclCommInitAll and one thread issues the per-device calls inside a group.
// NCCL is loaded at run time (dlopen of libnccl.so.2 — the copy PyTorch already mapped if there is one), so the
// library keeps loading on machines without NCCL; ckBroadcastUpload then uses its tree of peer copies instead.
#include "handles.h"

#include <dlfcn.h>

#include <map>
#include <mutex>
#include <string>
#include <vector>

namespace ck {

typedef struct ncclComm* ncclComm_t;
typedef int ncclResult_t;
typedef ncclResult_t (*CommInitAllFn)(ncclComm_t*, int, const int*);
typedef ncclResult_t (*GroupFn)(void);
typedef ncclResult_t (*BroadcastFn)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t);
typedef ncclResult_t (*CommDestroyFn)(ncclComm_t);
typedef const char* (*ErrStrFn)(ncclResult_t);

struct NcclApi {
    void* lib = nullptr;
    CommInitAllFn commInitAll = nullptr;
    GroupFn groupStart = nullptr, groupEnd = nullptr;
    BroadcastFn broadcast = nullptr;
    CommDestroyFn commDestroy = nullptr;
    ErrStrFn errStr = nullptr;
    bool ok = false;
};

static NcclApi& api() {
    static NcclApi a = [] {
        NcclApi x;
        const char* off = getenv("CK_BROADCAST");
        if (off && std::string(off) == "p2p") return x;
        for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
            x.lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (x.lib) break;
        }
        if (!x.lib) return x;
        x.commInitAll = (CommInitAllFn)dlsym(x.lib, "ncclCommInitAll");
        x.groupStart = (GroupFn)dlsym(x.lib, "ncclGroupStart");
        x.groupEnd = (GroupFn)dlsym(x.lib, "ncclGroupEnd");
        x.broadcast = (BroadcastFn)dlsym(x.lib, "ncclBroadcast");
        x.commDestroy = (CommDestroyFn)dlsym(x.lib, "ncclCommDestroy");
        x.errStr = (ErrStrFn)dlsym(x.lib, "ncclGetErrorString");
        x.ok = x.commInitAll && x.groupStart && x.groupEnd && x.broadcast && x.commDestroy;
        return x;
    }();
    return a;
}

static std::mutex g_nccl_mu;
static std::map<std::vector<int>, std::vector<ncclComm_t>> g_comms;  // one communicator set per ordered device list
static std::map<std::vector<int>, unsigned> g_root;                   // rotating root per device list

bool nccl_available() { return api().ok; }

// Broadcasts `bytes` bytes from ptrs[root] to every ptrs[i] (device i = ordinals[i], ordered on streams[i]).
// Returns 0 on success, <0 if NCCL is unavailable (caller falls back), >0 on an NCCL error.
int nccl_broadcast(const std::vector<int>& ordinals, const std::vector<void*>& ptrs, const std::vector<cudaStream_t>& streams,
                   size_t bytes, int root) {
    NcclApi& a = api();
    if (!a.ok) return -1;
    std::lock_guard<std::mutex> lk(g_nccl_mu);
    auto it = g_comms.find(ordinals);
    if (it == g_comms.end()) {
        std::vector<ncclComm_t> comms(ordinals.size(), nullptr);
        ncclResult_t r = a.commInitAll(comms.data(), (int)ordinals.size(), ordinals.data());
        if (r != 0) {
            set_error(std::string("ncclCommInitAll: ") + (a.errStr ? a.errStr(r) : "error"));
            return 1;
        }
        it = g_comms.emplace(ordinals, comms).first;
    }
    auto& comms = it->second;
    ncclResult_t r = a.groupStart();
    for (size_t i = 0; i < ordinals.size() && r == 0; i++) {
        cudaSetDevice(ordinals[i]);
        r = a.broadcast(ptrs[root], ptrs[i], bytes, /*ncclChar*/ 0, root, comms[i], streams[i]);
    }
    ncclResult_t e = a.groupEnd();
    if (r == 0) r = e;
    if (r != 0) {
        set_error(std::string("ncclBroadcast: ") + (a.errStr ? a.errStr(r) : "error"));
        return 2;
    }
    return 0;
}

unsigned nccl_next_root(const std::vector<int>& ordinals) {
    std::lock_guard<std::mutex> lk(g_nccl_mu);
    return g_root[ordinals]++ % (unsigned)ordinals.size();
}

}  // namespace ck
