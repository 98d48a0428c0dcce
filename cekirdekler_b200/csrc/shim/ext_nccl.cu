// ext_nccl.cu — NCCL broadcast over NVLink/NVSwitch for whole-array `read` inputs.
//
// The reference uploads a whole-`read` array from the host to EVERY device (Worker.cs:833-834, called per worker from
// Cores.cs:747-768).  Here the array crosses the host link once, into one device (the root rotates from call to call so
// that no single host link carries every upload), and NCCL broadcasts it from there to the other devices of the
// cruncher over NVLink.  One process drives all devices (the reference's execution model), so the communicators come
// from ncclCommInitAll and one thread issues the per-device calls inside a group.
// With one process per GPU (torchrun) the same broadcast runs over a communicator of the job's ranks instead
// (ckSpmdUniqueId / ckSpmdInit / ckSpmdBroadcast at the end of this file): rank 0 draws the id, the launcher's own channel
// carries it, every rank joins; the host layer then lets the root rank upload and all ranks broadcast in place.
// NCCL is loaded at run time (dlopen of libnccl.so.2 — the copy PyTorch already mapped if there is one), so the
// library keeps loading on machines without NCCL; ckBroadcastUpload then uses its tree of peer copies instead.
#include "handles.h"
#include "../../../include/KutuphaneCL.h"

#include <dlfcn.h>

#include <cstring>

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
struct ncclUniqueId { char internal[128]; };   // NCCL_UNIQUE_ID_BYTES
typedef ncclResult_t (*GetUniqueIdFn)(ncclUniqueId*);
typedef ncclResult_t (*CommInitRankFn)(ncclComm_t*, int, ncclUniqueId, int);

struct NcclApi {
    void* lib = nullptr;
    CommInitAllFn commInitAll = nullptr;
    GroupFn groupStart = nullptr, groupEnd = nullptr;
    BroadcastFn broadcast = nullptr;
    CommDestroyFn commDestroy = nullptr;
    ErrStrFn errStr = nullptr;
    GetUniqueIdFn getUniqueId = nullptr;
    CommInitRankFn commInitRank = nullptr;
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
        x.getUniqueId = (GetUniqueIdFn)dlsym(x.lib, "ncclGetUniqueId");
        x.commInitRank = (CommInitRankFn)dlsym(x.lib, "ncclCommInitRank");
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

// ---- one process per GPU: a communicator over the ranks of the job (ncclCommInitRank), for the same broadcast ----
static ncclComm_t g_rank_comm = nullptr;
static int g_rank = -1, g_world = 0;

}  // namespace ck

extern "C" {

// Rank 0 makes the 128-byte id and hands it to the other ranks by whatever channel the job has (torch.distributed here).
int32_t ckSpmdUniqueId(void* out128) {
    ck::NcclApi& a = ck::api();
    if (!a.ok || !a.getUniqueId || !a.commInitRank || !out128) {
        ck::set_error("ckSpmdUniqueId: NCCL is not available");
        return -1;
    }
    ck::ncclUniqueId id;
    const ck::ncclResult_t r = a.getUniqueId(&id);
    if (r != 0) {
        ck::set_error(std::string("ncclGetUniqueId: ") + (a.errStr ? a.errStr(r) : "error"));
        return 1;
    }
    memcpy(out128, id.internal, sizeof id.internal);
    return 0;
}

// Collective over all ranks of the job; `deviceOrdinal` is this rank's GPU.
int32_t ckSpmdInit(const void* id128, int32_t rank, int32_t world, int32_t deviceOrdinal) {
    ck::NcclApi& a = ck::api();
    if (!a.ok || !a.commInitRank || !id128 || world < 1 || rank < 0 || rank >= world) {
        ck::set_error("ckSpmdInit: NCCL is not available or bad arguments");
        return -1;
    }
    std::lock_guard<std::mutex> lk(ck::g_nccl_mu);
    if (ck::g_rank_comm) return (ck::g_rank == rank && ck::g_world == world) ? 0 : 2;
    if (!ck::cuda_ok(cudaSetDevice(deviceOrdinal), "ckSpmdInit: cudaSetDevice")) return 3;
    ck::ncclUniqueId id;
    memcpy(id.internal, id128, sizeof id.internal);
    const ck::ncclResult_t r = a.commInitRank(&ck::g_rank_comm, world, id, rank);
    if (r != 0) {
        ck::g_rank_comm = nullptr;
        ck::set_error(std::string("ncclCommInitRank: ") + (a.errStr ? a.errStr(r) : "error"));
        return 1;
    }
    ck::g_rank = rank;
    ck::g_world = world;
    return 0;
}

int32_t ckSpmdWorld(void) { return ck::g_rank_comm ? ck::g_world : 0; }

// In-place broadcast of the whole buffer from rank `root`, ordered on this rank's queue (every rank calls it).
int32_t ckSpmdBroadcast(ck_handle hQueue, ck_handle hBuffer, int32_t root) {
    ck::NcclApi& a = ck::api();
    if (!ck::g_rank_comm || !hQueue || !hBuffer || root < 0 || root >= ck::g_world) {
        ck::set_error("ckSpmdBroadcast: no communicator (ckSpmdInit) or bad arguments");
        return -1;
    }
    ck::Queue* q = ck::as<ck::Queue>(hQueue);
    ck::Buffer* b = ck::as<ck::Buffer>(hBuffer);
    if (b->zero_copy) {
        ck::set_error("ckSpmdBroadcast: zero-copy buffers alias host memory and are not broadcast");
        return -2;
    }
    cudaSetDevice(q->ordinal);
    const ck::ncclResult_t r = a.broadcast(b->dptr, b->dptr, (size_t)b->bytes(), /*ncclChar*/ 0, root, ck::g_rank_comm, q->stream);
    if (r != 0) {
        ck::set_error(std::string("ncclBroadcast: ") + (a.errStr ? a.errStr(r) : "error"));
        return 1;
    }
    return 0;
}

void ckSpmdFinalize(void) {
    std::lock_guard<std::mutex> lk(ck::g_nccl_mu);
    if (ck::g_rank_comm && ck::api().commDestroy) ck::api().commDestroy(ck::g_rank_comm);
    ck::g_rank_comm = nullptr;
    ck::g_rank = -1;
    ck::g_world = 0;
}

}  // extern "C"
