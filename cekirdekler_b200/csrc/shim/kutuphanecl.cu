// kutuphanecl.cu — libKutuphaneCL.so: the reference's 81-symbol native boundary on CUDA (B200, sm_100a).
//
// The reference's KutuphaneCL.dll is C++ over OpenCL and its source is not in the reference tree; what is
// in the tree is the C# side of the boundary (DllImport sites cited per function in include/KutuphaneCL.h).
// This file implements that boundary on the CUDA runtime:
//   platform      → one synthetic "NVIDIA CUDA" platform (0 CPUs, N GPUs, 0 accelerators)
//   context       → a device ordinal (primary context) with peer access enabled to every other GPU
//   command queue → one non-blocking cudaStream_t; markers are polled events
//   event / array → reference-counted cudaEvent_t / wait list (cudaStreamWaitEvent)
//   buffer        → cudaMalloc, or the mapped pinned host pointer itself for zero-copy ("stream") buffers
//   transfers     → cudaMemcpyAsync of element range [ref, ref+range) at the same offset on both sides
//   program       → NVRTC to a cubin, loaded as a context-independent cudaLibrary_t (program.cpp)
//   kernel        → cudaKernel_t + its own argument table (one object per name+computeId, Worker.cs:291-316)
//   compute       → cudaLaunchKernel(grid = global/local, block = local) with the device's start index
//                   passed as the hidden first parameter
//   host arrays   → cudaHostAlloc(portable|mapped), aligned; registered so createBuffer can alias them
// There is no CPU execution path: without a CUDA device every compute/transfer entry point fails loudly.
#include "handles.h"
#include "../../../include/KutuphaneCL.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>

namespace ck {

// ------------------------------------------------------------------ errors
static thread_local std::string t_last_error;
void set_error(const std::string& msg) {
    t_last_error = msg;
    if (getenv("CK_VERBOSE")) fprintf(stderr, "[KutuphaneCL] %s\n", msg.c_str());
}
bool cuda_ok(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return true;
    set_error(std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")");
    return false;
}
#define CK_CUDA(call) ck::cuda_ok((call), #call)

static std::atomic<uint64_t> g_launches{0};
static std::atomic<uint64_t> g_broadcasts_nccl{0}, g_broadcasts_p2p{0};

// ------------------------------------------------------------------ registry
Registry& Registry::get() {
    static Registry* r = new Registry();  // leaked on purpose: finalizer-thread deletes may come at exit
    return *r;
}
void Registry::add(Obj* o) {
    std::lock_guard<std::mutex> lk(mu_);
    live_.insert(o);
}
bool Registry::remove(void* o, Kind k) {
    if (!o) return false;
    std::lock_guard<std::mutex> lk(mu_);
    auto it = live_.find(o);
    if (it == live_.end()) return false;
    if (static_cast<Obj*>(o)->kind != k) return false;
    live_.erase(it);
    return true;
}
bool Registry::live(void* o, Kind k) {
    if (!o) return false;
    std::lock_guard<std::mutex> lk(mu_);
    auto it = live_.find(o);
    return it != live_.end() && static_cast<Obj*>(o)->kind == k;
}
template <typename T> static T* reg(T* o) {
    Registry::get().add(o);
    return o;
}
template <typename T> static void del(void* h, Kind k) {
    if (Registry::get().remove(h, k)) as<T>(h)->release();
}

// ------------------------------------------------------------------ device table
static int device_count() {
    static int n = [] {
        int c = 0;
        if (cudaGetDeviceCount(&c) != cudaSuccess) {
            cudaGetLastError();
            c = 0;
        }
        return c;
    }();
    return n;
}

static std::once_flag g_peer_once[64];
static void enable_peers(int ordinal) {
    // NVSwitch: every peer is reachable; enable once per device so kernels and copies can address peer memory
    if (ordinal < 0 || ordinal >= 64) return;
    std::call_once(g_peer_once[ordinal], [ordinal] {
        int n = device_count();
        cudaSetDevice(ordinal);
        for (int p = 0; p < n; p++) {
            if (p == ordinal) continue;
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, ordinal, p) == cudaSuccess && can) {
                cudaError_t e = cudaDeviceEnablePeerAccess(p, 0);
                if (e != cudaSuccess) cudaGetLastError();  // already enabled is fine
            }
        }
    });
}

// ------------------------------------------------------------------ object lifetimes
Platform::Platform() : Obj(Kind::Platform) {
    name = new String();
    name->s = "NVIDIA CUDA";
    vendor = new String();
    vendor->s = "NVIDIA Corporation";
}
Platform::~Platform() {
    name->release();
    vendor->release();
}

Event::~Event() {
    if (ev) {
        cudaSetDevice(ordinal);
        cudaEventDestroy(ev);
    }
}
bool Event::record(cudaStream_t s, int dev) {
    std::lock_guard<std::mutex> lk(mu);
    if (ev && ordinal != dev) {  // re-used on another device: re-create there
        cudaSetDevice(ordinal);
        cudaEventDestroy(ev);
        ev = nullptr;
        cudaSetDevice(dev);
    }
    if (!ev) {
        if (!CK_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming))) return false;
        ordinal = dev;
    }
    if (!CK_CUDA(cudaEventRecord(ev, s))) return false;
    recorded.store(true, std::memory_order_release);
    return true;
}

Queue::~Queue() {
    cudaSetDevice(ordinal);
    if (stream) {
        cudaStreamSynchronize(stream);
        cudaStreamDestroy(stream);
    }
    for (auto e : pending) cudaEventDestroy(e);
    for (auto e : pool) cudaEventDestroy(e);
    if (ctx) ctx->release();
}
void Queue::wait_list(EventArr* arr) {
    if (!arr) return;
    std::lock_guard<std::mutex> lk(arr->mu);
    for (Event* e : arr->evs) {
        if (!e->recorded.load(std::memory_order_acquire)) continue;  // never enqueued: nothing to wait for
        std::lock_guard<std::mutex> lk2(e->mu);
        if (e->ev) CK_CUDA(cudaStreamWaitEvent(stream, e->ev, 0));
    }
}

UserEvent::~UserEvent() {
    if (flag) cudaFreeHost((void*)flag);
    if (ctx) ctx->release();
}

Buffer::~Buffer() {
    if (owns && dptr) {
        cudaSetDevice(ordinal);
        cudaFree(dptr);
    }
    if (ctx) ctx->release();
}

Program::~Program() {
    if (lib) cudaLibraryUnload(lib);
    if (ctx) ctx->release();
}

Kernel::~Kernel() {
    for (auto* b : args)
        if (b) b->release();
    if (prog) prog->release();
}

Array::~Array() {
    if (raw) {
        pinned_remove(head);
        if (pinned) cudaFreeHost(raw);
        else free(raw);
    }
}

Timer::~Timer() {
    cudaSetDevice(ordinal);
    if (a) cudaEventDestroy(a);
    if (b) cudaEventDestroy(b);
}

// ------------------------------------------------------------------ pinned table
static std::mutex g_pin_mu;
static std::map<uintptr_t, int64_t> g_pinned;  // head → bytes
void pinned_add(void* p, int64_t bytes) {
    std::lock_guard<std::mutex> lk(g_pin_mu);
    g_pinned[(uintptr_t)p] = bytes;
}
void pinned_remove(void* p) {
    std::lock_guard<std::mutex> lk(g_pin_mu);
    g_pinned.erase((uintptr_t)p);
}
bool pinned_lookup(const void* p, int64_t* bytes_out) {
    std::lock_guard<std::mutex> lk(g_pin_mu);
    auto it = g_pinned.upper_bound((uintptr_t)p);
    if (it == g_pinned.begin()) return false;
    --it;
    if ((uintptr_t)p >= it->first + (uintptr_t)it->second) return false;
    if (bytes_out) *bytes_out = it->second - (int64_t)((uintptr_t)p - it->first);
    return true;
}

// ------------------------------------------------------------------ helpers
static int size_of_cl_type(int t) {
    // ClBuffer.SizeOf (ClBuffer.cs:266-275)
    switch (t) {
        case 0: return 4;  // cl_float
        case 1: return 8;  // cl_double
        case 2: return 4;  // cl_int
        case 3: return 8;  // cl_long
        case 4: return 2;  // cl_half  (C# char, UTF-16 code unit)
        case 5: return 1;  // cl_char  (C# byte)
        case 6: return 4;  // cl_uint
        default: return 0;
    }
}
static int size_of_arr_type(int t) {
    // CSpaceArrays.ARR_* (CSpaceArrays.cs:48-78)
    switch (t) {
        case 0: return 4;  // ARR_FLOAT
        case 1: return 8;  // ARR_DOUBLE
        case 2: return 8;  // ARR_LONG
        case 3: return 4;  // ARR_INT
        case 4: return 4;  // ARR_UINT
        case 5: return 1;  // ARR_BYTE
        case 6: return 2;  // ARR_CHAR
        default: return 0;
    }
}

enum class Dir { H2D, D2H };

static void transfer(void* hq, void* hb, int64_t ref, int64_t range, void* host, void* harr, void* hevt, Dir dir) {
    if (!hq || !hb) {
        set_error("transfer: null queue or buffer handle");
        return;
    }
    Queue* q = as<Queue>(hq);
    Buffer* b = as<Buffer>(hb);
    cudaSetDevice(q->ordinal);
    if (harr) q->wait_list(as<EventArr>(harr));
    if (range < 0) {  // whole buffer
        ref = 0;
        range = b->elems;
    }
    if (ref < 0 || ref + range > b->elems) {
        // the reference would overrun the buffer here; refuse and say so (the caller ignores the return value)
        set_error("transfer: element range [" + std::to_string(ref) + ", " + std::to_string(ref + range) + ") outside a buffer of " +
                  std::to_string(b->elems) + " elements — nothing copied");
        fprintf(stderr, "[KutuphaneCL] %s\n", t_last_error.c_str());
    } else if (!b->zero_copy && range > 0) {
        if (!host) {
            set_error("transfer: null host array");
        } else {
            const int64_t off = ref * (int64_t)b->elem_size;
            const int64_t bytes = range * (int64_t)b->elem_size;
            char* d = (char*)b->dptr + off;
            char* h = (char*)host + off;
            if (dir == Dir::H2D) CK_CUDA(cudaMemcpyAsync(d, h, (size_t)bytes, cudaMemcpyHostToDevice, q->stream));
            else CK_CUDA(cudaMemcpyAsync(h, d, (size_t)bytes, cudaMemcpyDeviceToHost, q->stream));
        }
    }
    if (hevt) as<Event>(hevt)->record(q->stream, q->ordinal);
}

static int launch(Queue* q, Kernel* k, int offset, int global, int local) {
    if (k->err || !k->fn) {
        set_error("compute: kernel '" + k->name + "' was not created successfully");
        return 1;
    }
    if (local <= 0 || global < 0 || (global % local) != 0) {
        set_error("compute: global range must be a non-negative multiple of the local range");
        return 2;
    }
    if (global == 0) return 0;
    if (local > 1024) {
        set_error("compute: local range above 1024 threads");
        return 3;
    }
    unsigned int goff = (unsigned int)offset;
    void* ptrs[64];
    void* params[65];
    int nargs;
    {
        std::lock_guard<std::mutex> lk(k->mu);
        nargs = (int)k->args.size();
        if (k->nparams >= 0 && nargs != k->nparams) {
            set_error("compute: kernel '" + k->name + "' takes " + std::to_string(k->nparams) + " buffer arguments but " +
                      std::to_string(nargs) + " are bound");
            return 4;
        }
        if (nargs > 64) {
            set_error("compute: more than 64 kernel arguments");
            return 5;
        }
        params[0] = &goff;
        for (int i = 0; i < nargs; i++) {
            if (!k->args[i]) {
                set_error("compute: kernel '" + k->name + "' argument " + std::to_string(i) + " is not bound");
                return 6;
            }
            ptrs[i] = k->args[i]->dptr;
            params[i + 1] = &ptrs[i];
        }
    }
    dim3 grid((unsigned)(global / local)), block((unsigned)local);
    cudaError_t e = cudaLaunchKernel((const void*)k->fn, grid, block, params, 0, q->stream);
    if (!cuda_ok(e, "cudaLaunchKernel")) return 100 + (int)e;
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return 0;
}

typedef int (*StreamWaitValue32Fn)(cudaStream_t, unsigned long long, uint32_t, unsigned int);
static StreamWaitValue32Fn stream_wait_value32() {
    static StreamWaitValue32Fn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult st;
        if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &p, cudaEnableDefault, &st) != cudaSuccess ||
            st != cudaDriverEntryPointSuccess) {
            cudaGetLastError();
            p = nullptr;
        }
        return (StreamWaitValue32Fn)p;
    }();
    return fn;
}

}  // namespace ck

using namespace ck;

extern "C" {

// ================================================================== platform list / platform
ck_handle platformList(void) { return reg(new PlatformList()); }
int32_t numberOfPlatforms(ck_handle) { return 1; }
void deletePlatformList(ck_handle h) { del<PlatformList>(h, Kind::PlatformList); }

ck_handle createPlatform(ck_handle, int32_t index) {
    if (index != 0) {
        set_error("createPlatform: only platform 0 (NVIDIA CUDA) exists");
        return nullptr;
    }
    return reg(new Platform());
}
void deletePlatform(ck_handle h) { del<Platform>(h, Kind::Platform); }
ck_handle getPlatformInfo(ck_handle h) { return h ? as<Platform>(h)->name : nullptr; }
ck_handle getPlatformNameString(ck_handle h) { return h ? as<Platform>(h)->name : nullptr; }
ck_handle getPlatformVendorNameString(ck_handle h) { return h ? as<Platform>(h)->vendor : nullptr; }
// OpenCL's CL_DEVICE_TYPE_* values; the C# treats them as opaque codes (ClPlatform.cs:62-72)
int32_t CODE_CPU(void) { return 2; }
int32_t CODE_GPU(void) { return 4; }
int32_t CODE_ACC(void) { return 8; }
int32_t numberOfCpusInPlatform(ck_handle) { return 0; }
int32_t numberOfGpusInPlatform(ck_handle) { return device_count(); }
int32_t numberOfAcceleratorsInPlatform(ck_handle) { return 0; }

// ================================================================== device
ck_handle createDevice(ck_handle hPlatform, int32_t type_, int32_t index_) {
    if (type_ != CODE_GPU()) {
        set_error("createDevice: this platform only has GPU devices");
        return nullptr;
    }
    if (index_ < 0 || index_ >= device_count()) {
        set_error("createDevice: no CUDA device with that index (is a GPU present?)");
        return nullptr;
    }
    cudaDeviceProp p;
    if (!CK_CUDA(cudaGetDeviceProperties(&p, index_))) return nullptr;
    Device* d = new Device();
    d->ordinal = index_;
    d->name = p.name;
    d->sms = p.multiProcessorCount;
    d->mem = (uint64_t)p.totalGlobalMem;
    d->cc_major = p.major;
    d->cc_minor = p.minor;
    if (hPlatform) {
        d->platform = as<Platform>(hPlatform);
        d->platform->retain();
    }
    return reg(d);
}
ck_handle createDeviceAsPartition(ck_handle hPlatform, int32_t type_, int32_t index_, int32_t) {
    // device fission only exists for OpenCL CPU devices (ClDevice.cs:85-95); a GPU is used whole
    return createDevice(hPlatform, type_, index_);
}
void deleteDevice(ck_handle h) { del<Device>(h, Kind::Device); }
void getDeviceName(ck_handle hDevice, ck_handle hString) {
    if (hDevice && hString) as<String>(hString)->s = as<Device>(hDevice)->name;
}
void getDeviceVendorName(ck_handle hDevice, ck_handle hString) {
    if (hDevice && hString) as<String>(hString)->s = "NVIDIA Corporation";
}
ck_bool deviceGDDR(ck_handle) { return 1; }  // dedicated memory (HBM3e)
int32_t deviceComputeUnits(ck_handle h) { return h ? as<Device>(h)->sms : 0; }
uint64_t deviceMemSize(ck_handle h) { return h ? as<Device>(h)->mem : 0; }

// ================================================================== context
ck_handle createContext(ck_handle, ck_handle hDevice) {
    if (!hDevice) {
        set_error("createContext: null device");
        return nullptr;
    }
    Device* d = as<Device>(hDevice);
    if (!CK_CUDA(cudaSetDevice(d->ordinal))) return nullptr;
    if (!CK_CUDA(cudaFree(0))) return nullptr;  // force primary-context creation
    enable_peers(d->ordinal);
    Context* c = new Context();
    c->ordinal = d->ordinal;
    c->device = d;
    d->retain();
    return reg(c);
}
void deleteContext(ck_handle h) { del<Context>(h, Kind::Context); }

// ================================================================== command queue
ck_handle createCommandQueue(ck_handle hContext, ck_handle, int32_t) {
    if (!hContext) {
        set_error("createCommandQueue: null context");
        return nullptr;
    }
    Context* c = as<Context>(hContext);
    if (!CK_CUDA(cudaSetDevice(c->ordinal))) return nullptr;
    cudaStream_t s;
    if (!CK_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking))) return nullptr;
    Queue* q = new Queue();
    q->ctx = c;
    c->retain();
    q->ordinal = c->ordinal;
    q->stream = s;
    return reg(q);
}
ck_handle createCommandQueue2(ck_handle hContext, ck_handle hDevice, int32_t async_, ck_bool) {
    // the OpenCL 2.0 "default device queue" (Worker.cs:224-227) has no CUDA counterpart: device-side launches
    // need no queue object.  An ordinary queue is returned so that the handle stays valid.
    return createCommandQueue(hContext, hDevice, async_);
}
void deleteCommandQueue(ck_handle h) { del<Queue>(h, Kind::Queue); }

void addMarkerToCommandQueue(ck_handle h) {
    if (!h) return;
    Queue* q = as<Queue>(h);
    cudaSetDevice(q->ordinal);
    std::lock_guard<std::mutex> lk(q->mu);
    cudaEvent_t e;
    if (!q->pool.empty()) {
        e = q->pool.back();
        q->pool.pop_back();
    } else if (!CK_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming))) {
        return;
    }
    CK_CUDA(cudaEventRecord(e, q->stream));
    q->pending.push_back(e);
}
int32_t getMarkerCounterOfCommandQueue(ck_handle h) {
    if (!h) return 0;
    Queue* q = as<Queue>(h);
    std::lock_guard<std::mutex> lk(q->mu);
    while (!q->pending.empty()) {
        cudaError_t e = cudaEventQuery(q->pending.front());
        if (e == cudaErrorNotReady) break;
        if (e != cudaSuccess) cudaGetLastError();
        q->pool.push_back(q->pending.front());
        q->pending.pop_front();
        q->reached++;
    }
    return q->reached;
}
void resetMarkerCounterOfCommandQueue(ck_handle h) {
    if (!h) return;
    Queue* q = as<Queue>(h);
    std::lock_guard<std::mutex> lk(q->mu);
    q->reached = 0;
}

// ================================================================== events
ck_handle createEvent(void) { return reg(new Event()); }
void deleteEvent(ck_handle h) { del<Event>(h, Kind::Event); }
ck_handle createEventArr(ck_bool) { return reg(new EventArr()); }
void addToEventArr(ck_handle hArr, ck_handle hEvent, ck_bool) {
    if (!hArr || !hEvent) return;
    EventArr* a = as<EventArr>(hArr);
    Event* e = as<Event>(hEvent);
    e->retain();
    std::lock_guard<std::mutex> lk(a->mu);
    a->evs.push_back(e);
}
void deleteEventArr(ck_handle h) { del<EventArr>(h, Kind::EventArr); }

ck_handle createUserEvent(ck_handle hContext) {
    if (!hContext) return nullptr;
    Context* c = as<Context>(hContext);
    cudaSetDevice(c->ordinal);
    UserEvent* u = new UserEvent();
    void* p = nullptr;
    if (!CK_CUDA(cudaHostAlloc(&p, 64, cudaHostAllocPortable | cudaHostAllocMapped))) {
        delete u;
        return nullptr;
    }
    u->flag = (volatile int*)p;
    *u->flag = 0;
    u->ctx = c;
    c->retain();
    return reg(u);
}
void deleteUserEvent(ck_handle h) {
    if (Registry::get().live(h, Kind::UserEvent)) triggerUserEvent(h);  // never leave a queue gated forever
    del<UserEvent>(h, Kind::UserEvent);
}
void triggerUserEvent(ck_handle h) {
    if (!h) return;
    UserEvent* u = as<UserEvent>(h);
    std::lock_guard<std::mutex> lk(u->mu);
    __atomic_store_n((int*)u->flag, 1, __ATOMIC_RELEASE);
}
void addUserEvent(ck_handle hQueue, ck_handle hUserEvent) {
    if (!hQueue || !hUserEvent) return;
    Queue* q = as<Queue>(hQueue);
    UserEvent* u = as<UserEvent>(hUserEvent);
    auto fn = stream_wait_value32();
    if (!fn) {
        set_error("addUserEvent: stream memory-wait is not available in this driver");
        return;
    }
    cudaSetDevice(q->ordinal);
    void* d = nullptr;
    if (!CK_CUDA(cudaHostGetDevicePointer(&d, (void*)u->flag, 0))) return;
    int r = fn(q->stream, (unsigned long long)(uintptr_t)d, 1u, 0x0 /* CU_STREAM_WAIT_VALUE_GEQ: flag >= 1 */);
    if (r != 0) set_error("addUserEvent: cuStreamWaitValue32 failed with code " + std::to_string(r));
}
void incrementUserEvent(ck_handle h) {
    if (!h) return;
    UserEvent* u = as<UserEvent>(h);
    std::lock_guard<std::mutex> lk(u->mu);
    u->counter++;
}
void decrementUserEvent(ck_handle h, ck_handle) {
    if (!h) return;
    UserEvent* u = as<UserEvent>(h);
    bool fire;
    {
        std::lock_guard<std::mutex> lk(u->mu);
        fire = (--u->counter <= 0);
    }
    if (fire) triggerUserEvent(h);
}

// ================================================================== strings
ck_handle createString(void) { return reg(new String()); }
ck_handle writeToString(ck_handle h, const char* text) {
    if (h) as<String>(h)->s = text ? text : "";
    return h;
}
void deleteString(ck_handle h) { del<String>(h, Kind::String); }
const char* readFromString(ck_handle h) { return h ? as<String>(h)->s.c_str() : ""; }
void jsonStringCallBack(ck_handle h) { del<String>(h, Kind::String); }  // no-op for platform-owned strings

// ================================================================== program / kernel
ck_handle createProgram(ck_handle hContext, ck_handle hDevice, ck_handle hString) {
    Program* p = new Program();
    reg(p);
    if (!hContext || !hDevice || !hString) {
        p->err = 1;
        p->log = "createProgram: null handle";
        return p;
    }
    Context* c = as<Context>(hContext);
    Device* d = as<Device>(hDevice);
    p->ctx = c;
    c->retain();
    if (d->cc_major < 10) {
        p->err = 2;
        p->log = "createProgram: this library targets Blackwell (sm_100a); device '" + d->name + "' is sm_" +
                 std::to_string(d->cc_major) + std::to_string(d->cc_minor);
        set_error(p->log);
        return p;
    }
    p->cubin = compile_source(as<String>(hString)->s, d->cc_major, d->cc_minor);
    p->err = p->cubin->err;
    p->log = p->cubin->log;
    if (p->err) {
        set_error(p->log);
        return p;
    }
    cudaSetDevice(c->ordinal);
    cudaError_t e = cudaLibraryLoadData(&p->lib, p->cubin->image.data(), nullptr, nullptr, 0, nullptr, nullptr, 0);
    if (!cuda_ok(e, "cudaLibraryLoadData")) {
        p->err = 2000 + (int)e;
        p->log += std::string("\ncudaLibraryLoadData: ") + cudaGetErrorString(e);
        p->lib = nullptr;
    }
    return p;
}
void deleteProgram(ck_handle h) { del<Program>(h, Kind::Program); }
int32_t getProgramErr(ck_handle h) { return h ? as<Program>(h)->err : 1; }
char* readProgramErrString(ck_handle h) {
    // declared `string` on the C# side (ClProgram.cs:40): the CLR frees the returned pointer, so hand out a copy
    const std::string& s = h ? as<Program>(h)->log : std::string("null program");
    char* out = (char*)malloc(s.size() + 1);
    memcpy(out, s.c_str(), s.size() + 1);
    return out;
}

ck_handle createKernel(ck_handle hProgram, ck_handle hString) {
    Kernel* k = new Kernel();
    reg(k);
    if (!hProgram || !hString) {
        k->err = 1;
        return k;
    }
    Program* p = as<Program>(hProgram);
    k->prog = p;
    p->retain();
    k->name = as<String>(hString)->s;
    if (p->err || !p->lib) {
        k->err = 2;
        set_error("createKernel: program did not build");
        return k;
    }
    cudaSetDevice(p->ctx->ordinal);
    cudaError_t e = cudaLibraryGetKernel(&k->fn, p->lib, k->name.c_str());
    if (!cuda_ok(e, "cudaLibraryGetKernel")) {
        k->err = 3000 + (int)e;
        k->fn = nullptr;
        return k;
    }
    for (auto& kv : p->cubin->kernels)
        if (kv.first == k->name) k->nparams = kv.second;
    // Load the kernel's code onto the device NOW.  CUDA loads a library kernel lazily at its first launch, and that load
    // waits for the whole device; a first launch enqueued behind a user event that has not fired yet (ClUserEvent.cs:30-46:
    // the host gates a queue, fills it, then triggers) would then wait for the gate inside cudaLaunchKernel and never return
    // to the thread that is about to open it (seen on B200, round 2).  Asking for the attributes forces the load.
    cudaFuncAttributes attr;
    if (cudaFuncGetAttributes(&attr, (const void*)k->fn) != cudaSuccess) cudaGetLastError();
    return k;
}
void deleteKernel(ck_handle h) { del<Kernel>(h, Kind::Kernel); }
int32_t getKernelErr(ck_handle h) { return h ? as<Kernel>(h)->err : 1; }
void setKernelArgument(ck_handle hKernel, ck_handle hBuffer, int32_t index) {
    if (!hKernel || !hBuffer || index < 0 || index >= 64) {
        set_error("setKernelArgument: bad handle or index");
        return;
    }
    Kernel* k = as<Kernel>(hKernel);
    Buffer* b = as<Buffer>(hBuffer);
    b->retain();
    std::lock_guard<std::mutex> lk(k->mu);
    if ((int)k->args.size() <= index) k->args.resize(index + 1, nullptr);
    if (k->args[index]) k->args[index]->release();
    k->args[index] = b;
}
void setKernelArguments(ck_handle hKernel, ck_handle hBuffer, int32_t index_) { setKernelArgument(hKernel, hBuffer, index_); }

ck_handle createNdRange(int32_t n) { return reg(new NdRange(n)); }
void deleteNdRange(ck_handle h) { del<NdRange>(h, Kind::NdRange); }

// ================================================================== launch
int32_t compute(ck_handle hq, ck_handle hk, ck_handle hoff, ck_handle hglob, ck_handle hloc) {
    if (!hq || !hk || !hoff || !hglob || !hloc) {
        set_error("compute: null handle");
        return 1;
    }
    Queue* q = as<Queue>(hq);
    cudaSetDevice(q->ordinal);
    return launch(q, as<Kernel>(hk), as<NdRange>(hoff)->n, as<NdRange>(hglob)->n, as<NdRange>(hloc)->n);
}
int32_t computeRepeated(ck_handle hq, ck_handle hk, ck_handle hoff, ck_handle hglob, ck_handle hloc, int32_t repeats) {
    int err = 0;
    for (int i = 0; i < repeats && !err; i++) err = compute(hq, hk, hoff, hglob, hloc);
    return err;
}
int32_t computeRepeatedWithSyncKernel(ck_handle hq, ck_handle hk, ck_handle hoff, ck_handle hglob, ck_handle hloc,
                                      int32_t repeats, ck_handle hSyncKernel, ck_handle hZeroRange) {
    // the sync kernel runs one work-group at offset 0 between repeats (Cores.cs:793, Worker.cs:1088-1094)
    int err = 0;
    for (int i = 0; i < repeats && !err; i++) {
        err = compute(hq, hk, hoff, hglob, hloc);
        if (!err) err = compute(hq, hSyncKernel, hZeroRange, hloc, hloc);
    }
    return err;
}
int32_t computeEvent(ck_handle hq, ck_handle hk, ck_handle hoff, ck_handle hglob, ck_handle hloc, ck_handle hEventArr,
                     ck_handle hEvt) {
    if (!hq || !hk || !hoff || !hglob || !hloc) {
        set_error("computeEvent: null handle");
        return 1;
    }
    Queue* q = as<Queue>(hq);
    cudaSetDevice(q->ordinal);
    if (hEventArr) q->wait_list(as<EventArr>(hEventArr));
    int err = launch(q, as<Kernel>(hk), as<NdRange>(hoff)->n, as<NdRange>(hglob)->n, as<NdRange>(hloc)->n);
    if (hEvt) as<Event>(hEvt)->record(q->stream, q->ordinal);
    return err;
}

// ================================================================== synchronisation
void finish(ck_handle h) {
    if (!h) return;
    Queue* q = as<Queue>(h);
    cudaSetDevice(q->ordinal);
    CK_CUDA(cudaStreamSynchronize(q->stream));
}
void flush(ck_handle h) {
    if (!h) return;
    Queue* q = as<Queue>(h);
    cudaSetDevice(q->ordinal);
    if (cudaStreamQuery(q->stream) != cudaSuccess) cudaGetLastError();  // kicks submission; not-ready is expected
}
void wait2(ck_handle a, ck_handle b) {
    finish(a);
    finish(b);
}
void wait3(ck_handle a, ck_handle b, ck_handle c) {
    finish(a);
    finish(b);
    finish(c);
}
void waitN(ck_handle* list, ck_handle toSync, int32_t n) {
    // Worker.cs:364-420: every queue used in enqueue mode is drained, then the main queue
    for (int i = 0; list && i < n; i++) finish(list[i]);
    finish(toSync);
}

// ================================================================== buffers and transfers
ck_handle createBuffer(ck_handle hContext, int32_t numberOfElements, int32_t clArrayType, int32_t, void* arrayPointer,
                       ck_bool GDDR_BUFFER, ck_bool, ck_bool) {
    if (!hContext) {
        set_error("createBuffer: null context");
        return nullptr;
    }
    int es = size_of_cl_type(clArrayType);
    if (!es || numberOfElements < 0) {
        set_error("createBuffer: bad element type or count");
        return nullptr;
    }
    Context* c = as<Context>(hContext);
    if (!CK_CUDA(cudaSetDevice(c->ordinal))) return nullptr;
    Buffer* b = new Buffer();
    b->ordinal = c->ordinal;
    b->elems = numberOfElements;
    b->elem_size = es;
    int64_t avail = 0;
    if (!GDDR_BUFFER && arrayPointer && pinned_lookup(arrayPointer, &avail) && avail >= b->bytes()) {
        // "stream" buffer (Worker.cs:578, :611): kernels dereference the mapped pinned host array directly
        void* d = nullptr;
        if (CK_CUDA(cudaHostGetDevicePointer(&d, arrayPointer, 0))) {
            b->dptr = d;
            b->zero_copy = true;
            b->owns = false;
        }
    }
    if (!b->dptr) {
        size_t bytes = (size_t)b->bytes();
        if (bytes == 0) bytes = 16;
        if (!CK_CUDA(cudaMalloc(&b->dptr, bytes))) {
            delete b;
            return nullptr;
        }
        // device-resident arrays (no read/write flag) are never uploaded: start them at zero like the host arrays they mirror
        // (completed before the handle is returned: the queues are non-blocking streams, so a clear still pending on the
        // default stream would otherwise be free to land after the first kernel has written the buffer)
        if (CK_CUDA(cudaMemsetAsync(b->dptr, 0, bytes, cudaStreamPerThread))) CK_CUDA(cudaStreamSynchronize(cudaStreamPerThread));
    }
    b->ctx = c;
    c->retain();
    return reg(b);
}
void deleteBuffer(ck_handle h) { del<Buffer>(h, Kind::Buffer); }

ck_handle writeToBuffer(ck_handle q, ck_handle b, void* host) {
    transfer(q, b, 0, -1, host, nullptr, nullptr, Dir::H2D);
    return nullptr;
}
ck_handle writeToBufferEvent(ck_handle q, ck_handle b, void* host, ck_handle arr, ck_handle evt) {
    transfer(q, b, 0, -1, host, arr, evt, Dir::H2D);
    return nullptr;
}
ck_handle writeToBufferRanged(ck_handle q, ck_handle b, int32_t ref, int32_t range, void* host) {
    transfer(q, b, ref, range < 0 ? 0 : range, host, nullptr, nullptr, Dir::H2D);
    return nullptr;
}
ck_handle writeToBufferRangedEvent(ck_handle q, ck_handle b, int32_t ref, int32_t range, void* host, ck_handle arr,
                                   ck_handle evt) {
    transfer(q, b, ref, range < 0 ? 0 : range, host, arr, evt, Dir::H2D);
    return nullptr;
}
ck_handle readFromBuffer(ck_handle q, ck_handle b, void* host) {
    transfer(q, b, 0, -1, host, nullptr, nullptr, Dir::D2H);
    return nullptr;
}
ck_handle readFromBufferEvent(ck_handle q, ck_handle b, void* host, ck_handle arr, ck_handle evt) {
    transfer(q, b, 0, -1, host, arr, evt, Dir::D2H);
    return nullptr;
}
ck_handle readFromBufferRanged(ck_handle q, ck_handle b, int32_t ref, int32_t range, void* host) {
    transfer(q, b, ref, range < 0 ? 0 : range, host, nullptr, nullptr, Dir::D2H);
    return nullptr;
}
ck_handle readFromBufferRangedEvent(ck_handle q, ck_handle b, int32_t ref, int32_t range, void* host, ck_handle arr,
                                    ck_handle evt) {
    transfer(q, b, ref, range < 0 ? 0 : range, host, arr, evt, Dir::D2H);
    return nullptr;
}

// ================================================================== host arrays
void sizeOf(int32_t* t) {
    if (!t) return;
    for (int i = 0; i < 20; i++) t[i] = size_of_arr_type(i);
}
ck_handle createArray(int32_t numberOfElements, int32_t alignment, int32_t typeOfArray) {
    int es = size_of_arr_type(typeOfArray);
    if (!es || numberOfElements < 0) {
        set_error("createArray: bad element type or count");
        return nullptr;
    }
    if (alignment < 64) alignment = 64;
    int64_t payload = (int64_t)numberOfElements * es;
    int64_t total = payload + alignment + 64;
    Array* a = new Array();
    void* raw = nullptr;
    if (device_count() > 0 && cudaHostAlloc(&raw, (size_t)total, cudaHostAllocPortable | cudaHostAllocMapped) == cudaSuccess) {
        a->pinned = true;
    } else {
        // no CUDA device in this process (CPU-only test box): plain memory; nothing can be computed with it
        cudaGetLastError();
        raw = malloc((size_t)total);
        if (!raw) {
            delete a;
            set_error("createArray: out of host memory");
            return nullptr;
        }
    }
    a->raw = raw;
    uintptr_t p = (uintptr_t)raw;
    p = (p + (uintptr_t)alignment - 1) / (uintptr_t)alignment * (uintptr_t)alignment;
    a->head = (void*)p;
    a->bytes = payload;
    memset(a->head, 0, (size_t)payload);
    if (a->pinned) pinned_add(a->head, payload > 0 ? payload : 1);
    return reg(a);
}
void* alignedArrHead(ck_handle h) { return h ? as<Array>(h)->head : nullptr; }
void deleteArray(ck_handle h) { del<Array>(h, Kind::Array); }
void copyMemory(void* dest, void* src, uint32_t count) {
    if (dest && src && count) memmove(dest, src, count);
}

// ================================================================== extensions
const char* ckGetLastError(void) { return t_last_error.c_str(); }
uint64_t ckLaunchCount(void) { return g_launches.load(); }
int32_t ckDeviceCount(void) { return device_count(); }

ck_handle ckCreateTimer(ck_handle hContext) {
    if (!hContext) return nullptr;
    Context* c = as<Context>(hContext);
    if (!CK_CUDA(cudaSetDevice(c->ordinal))) return nullptr;
    Timer* t = new Timer();
    t->ordinal = c->ordinal;
    if (!CK_CUDA(cudaEventCreate(&t->a)) || !CK_CUDA(cudaEventCreate(&t->b))) {
        delete t;
        return nullptr;
    }
    return reg(t);
}
void ckDeleteTimer(ck_handle h) { del<Timer>(h, Kind::Timer); }
void ckTimerStart(ck_handle h, ck_handle hq) {
    if (!h || !hq) return;
    Timer* t = as<Timer>(h);
    Queue* q = as<Queue>(hq);
    cudaSetDevice(q->ordinal);
    t->started = CK_CUDA(cudaEventRecord(t->a, q->stream));
    t->stopped = false;
}
void ckTimerStop(ck_handle h, ck_handle hq) {
    if (!h || !hq) return;
    Timer* t = as<Timer>(h);
    Queue* q = as<Queue>(hq);
    cudaSetDevice(q->ordinal);
    t->stopped = CK_CUDA(cudaEventRecord(t->b, q->stream));
}
double ckTimerElapsedMs(ck_handle h) {
    if (!h) return -1.0;
    Timer* t = as<Timer>(h);
    if (!t->started || !t->stopped) return -1.0;
    cudaSetDevice(t->ordinal);
    if (!CK_CUDA(cudaEventSynchronize(t->b))) return -1.0;
    float ms = 0;
    if (!CK_CUDA(cudaEventElapsedTime(&ms, t->a, t->b))) return -1.0;
    return (double)ms;
}

void* ckBufferDevicePointer(ck_handle h) { return h ? as<Buffer>(h)->dptr : nullptr; }
int64_t ckBufferBytes(ck_handle h) { return h ? as<Buffer>(h)->bytes() : 0; }
int32_t ckBufferIsZeroCopy(ck_handle h) { return h ? (as<Buffer>(h)->zero_copy ? 1 : 0) : 0; }
void* ckQueueStream(ck_handle h) { return h ? (void*)as<Queue>(h)->stream : nullptr; }
int32_t ckContextDeviceOrdinal(ck_handle h) { return h ? as<Context>(h)->ordinal : -1; }

int32_t ckPeerAccess(int32_t a, int32_t b) {
    int can = 0;
    if (a == b) return 1;
    if (cudaDeviceCanAccessPeer(&can, a, b) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return can;
}

int32_t ckCopyBufferPeer(ck_handle hq, ck_handle hDst, ck_handle hSrc, int32_t ref, int32_t range, ck_handle hArr,
                         ck_handle hEvt) {
    if (!hq || !hDst || !hSrc) {
        set_error("ckCopyBufferPeer: null handle");
        return 1;
    }
    Queue* q = as<Queue>(hq);
    Buffer* d = as<Buffer>(hDst);
    Buffer* s = as<Buffer>(hSrc);
    cudaSetDevice(q->ordinal);
    if (hArr) q->wait_list(as<EventArr>(hArr));
    int rc = 0;
    if (d->elem_size != s->elem_size || ref < 0 || range < 0 || (int64_t)ref + range > d->elems ||
        (int64_t)ref + range > s->elems) {
        set_error("ckCopyBufferPeer: element range or type mismatch");
        rc = 2;
    } else if (range > 0) {
        int64_t off = (int64_t)ref * d->elem_size, bytes = (int64_t)range * d->elem_size;
        cudaError_t e;
        if (d->ordinal == s->ordinal || d->zero_copy || s->zero_copy)
            e = cudaMemcpyAsync((char*)d->dptr + off, (char*)s->dptr + off, (size_t)bytes, cudaMemcpyDefault, q->stream);
        else
            e = cudaMemcpyPeerAsync((char*)d->dptr + off, d->ordinal, (char*)s->dptr + off, s->ordinal, (size_t)bytes,
                                    q->stream);
        if (!cuda_ok(e, "ckCopyBufferPeer")) rc = 100 + (int)e;
    }
    if (hEvt) as<Event>(hEvt)->record(q->stream, q->ordinal);
    return rc;
}

void ckQueueWaitEvent(ck_handle hq, ck_handle hEvent) {
    if (!hq || !hEvent) return;
    Queue* q = as<Queue>(hq);
    Event* e = as<Event>(hEvent);
    if (!e->recorded.load(std::memory_order_acquire)) return;
    cudaSetDevice(q->ordinal);
    std::lock_guard<std::mutex> lk(e->mu);
    if (e->ev) CK_CUDA(cudaStreamWaitEvent(q->stream, e->ev, 0));
}
void ckQueueRecordEvent(ck_handle hq, ck_handle hEvent) {
    if (!hq || !hEvent) return;
    Queue* q = as<Queue>(hq);
    cudaSetDevice(q->ordinal);
    as<Event>(hEvent)->record(q->stream, q->ordinal);
}
void ckEventSynchronize(ck_handle hEvent) {
    if (!hEvent) return;
    Event* e = as<Event>(hEvent);
    if (!e->recorded.load(std::memory_order_acquire)) return;
    cudaEvent_t ev;
    {
        std::lock_guard<std::mutex> lk(e->mu);
        ev = e->ev;
    }
    if (ev) CK_CUDA(cudaEventSynchronize(ev));
}

int32_t ckHostRegister(void* ptr, uint64_t bytes) {
    if (!ptr || !bytes) return 1;
    if (device_count() <= 0) {
        set_error("ckHostRegister: no CUDA device");
        return 2;
    }
    cudaError_t e = cudaHostRegister(ptr, (size_t)bytes, cudaHostRegisterPortable | cudaHostRegisterMapped);
    if (e == cudaErrorHostMemoryAlreadyRegistered) {
        cudaGetLastError();
        return 0;
    }
    if (!cuda_ok(e, "cudaHostRegister")) return 100 + (int)e;
    pinned_add(ptr, (int64_t)bytes);
    return 0;
}
int32_t ckHostUnregister(void* ptr) {
    if (!ptr) return 1;
    pinned_remove(ptr);
    cudaError_t e = cudaHostUnregister(ptr);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return 100 + (int)e;
    }
    return 0;
}

int32_t ckZeroBuffer(ck_handle hq, ck_handle hb) {
    if (!hq || !hb) return 1;
    Queue* q = as<Queue>(hq);
    Buffer* b = as<Buffer>(hb);
    cudaSetDevice(q->ordinal);
    return CK_CUDA(cudaMemsetAsync(b->dptr, 0, (size_t)b->bytes(), q->stream)) ? 0 : 2;
}

void ckBroadcastCounts(uint64_t* viaNccl, uint64_t* viaPeerCopies) {
    if (viaNccl) *viaNccl = g_broadcasts_nccl.load();
    if (viaPeerCopies) *viaPeerCopies = g_broadcasts_p2p.load();
}

int32_t ckBroadcastUpload(ck_handle* hQueues, ck_handle* hBuffers, int32_t n, void* hostArray) {
    if (!hQueues || !hBuffers || n <= 0 || !hostArray) {
        set_error("ckBroadcastUpload: bad arguments");
        return 1;
    }
    Buffer* b0 = as<Buffer>(hBuffers[0]);
    const int64_t bytes = b0->bytes();
    for (int i = 1; i < n; i++)
        if (as<Buffer>(hBuffers[i])->bytes() != bytes) {
            set_error("ckBroadcastUpload: buffers differ in size");
            return 2;
        }
    // Preferred path: one host-link upload into a rotating root, then ncclBroadcast over NVLink (ext_nccl.cu).
    bool anyZeroCopy = false, distinct = true;
    std::vector<int> ords;
    std::vector<void*> ptrs;
    std::vector<cudaStream_t> streams;
    for (int i = 0; i < n; i++) {
        Buffer* b = as<Buffer>(hBuffers[i]);
        anyZeroCopy = anyZeroCopy || b->zero_copy;
        for (int o : ords) distinct = distinct && (o != b->ordinal);
        ords.push_back(b->ordinal);
        ptrs.push_back(b->dptr);
        streams.push_back(as<Queue>(hQueues[i])->stream);
    }
    if (n > 1 && !anyZeroCopy && distinct && bytes > 0 && nccl_available()) {
        const int root = (int)nccl_next_root(ords);
        Queue* qr = as<Queue>(hQueues[root]);
        cudaSetDevice(qr->ordinal);
        if (CK_CUDA(cudaMemcpyAsync(ptrs[root], hostArray, (size_t)bytes, cudaMemcpyHostToDevice, qr->stream))) {
            const int r = nccl_broadcast(ords, ptrs, streams, (size_t)bytes, root);
            if (r == 0) {
                g_broadcasts_nccl.fetch_add(1, std::memory_order_relaxed);
                return 0;
            }
            if (r > 0 && getenv("CK_VERBOSE")) fprintf(stderr, "[KutuphaneCL] NCCL broadcast failed (%s); using peer copies\n", t_last_error.c_str());
        }
    }
    g_broadcasts_p2p.fetch_add(1, std::memory_order_relaxed);
    // Otherwise: one host-link upload, then a binary tree of NVLink peer copies: have[] = event marking "data present"
    std::vector<cudaEvent_t> have(n, nullptr);
    int rc = 0;
    auto mark = [&](int i) {
        Queue* q = as<Queue>(hQueues[i]);
        cudaSetDevice(q->ordinal);
        if (!CK_CUDA(cudaEventCreateWithFlags(&have[i], cudaEventDisableTiming))) rc = 3;
        else if (!CK_CUDA(cudaEventRecord(have[i], q->stream))) rc = 3;
    };
    {
        Queue* q0 = as<Queue>(hQueues[0]);
        cudaSetDevice(q0->ordinal);
        if (!b0->zero_copy && bytes > 0 &&
            !CK_CUDA(cudaMemcpyAsync(b0->dptr, hostArray, (size_t)bytes, cudaMemcpyHostToDevice, q0->stream)))
            rc = 4;
        mark(0);
    }
    for (int span = 1; span < n && !rc; span *= 2) {
        for (int src = 0; src < span && src + span < n; src++) {
            int dst = src + span;
            Queue* qd = as<Queue>(hQueues[dst]);
            Buffer* bs = as<Buffer>(hBuffers[src]);
            Buffer* bd = as<Buffer>(hBuffers[dst]);
            cudaSetDevice(qd->ordinal);
            CK_CUDA(cudaStreamWaitEvent(qd->stream, have[src], 0));
            if (!bd->zero_copy && bytes > 0) {
                cudaError_t e = (bs->ordinal == bd->ordinal)
                                    ? cudaMemcpyAsync(bd->dptr, bs->dptr, (size_t)bytes, cudaMemcpyDefault, qd->stream)
                                    : cudaMemcpyPeerAsync(bd->dptr, bd->ordinal, bs->dptr, bs->ordinal, (size_t)bytes,
                                                          qd->stream);
                if (!cuda_ok(e, "ckBroadcastUpload peer copy")) rc = 5;
            }
            mark(dst);
        }
    }
    for (int i = 0; i < n; i++)
        if (have[i]) {
            cudaSetDevice(as<Queue>(hQueues[i])->ordinal);
            cudaEventDestroy(have[i]);  // deferred by the runtime until the recorded work completes
        }
    return rc;
}

}  // extern "C"
