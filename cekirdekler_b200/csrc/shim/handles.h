// handles.h — object model behind the opaque handles of libKutuphaneCL.so.
//
// The reference's C# wrappers (ClDevice.cs, ClContext.cs, ClCommandQueue.cs, ClBuffer.cs, ClKernel.cs,
// ClProgram.cs, ClEvent.cs, ClEventArray.cs, ...) each hold one IntPtr from createX and call deleteX once,
// possibly from the GC finalizer thread and in any order (SURVEY.md §8b "Ownership").  So every object here
// is intrusively reference counted, children retain their parents, and a process-wide registry of live
// handles turns a stale or repeated delete into a no-op instead of a crash.
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <cstdint>
#include <deque>
#include <memory>
#include <mutex>
#include <string>
#include <unordered_set>
#include <vector>

namespace ck {

enum class Kind : uint32_t {
    PlatformList = 0x434b0001, Platform, Device, Context, Queue, Event, EventArr, UserEvent, String, Program,
    Kernel, NdRange, Buffer, Array, Timer
};

struct Obj {
    const Kind kind;
    std::atomic<int> rc{1};
    explicit Obj(Kind k) : kind(k) {}
    virtual ~Obj() {}
    void retain() { rc.fetch_add(1, std::memory_order_relaxed); }
    void release() {
        if (rc.fetch_sub(1, std::memory_order_acq_rel) == 1) delete this;
    }
};

// live-handle registry: only touched by create*/delete* (never on the transfer/launch path)
class Registry {
  public:
    static Registry& get();
    void add(Obj* o);
    // removes `o` if it is live and of kind `k`; returns whether the caller now owns the handle's reference
    bool remove(void* o, Kind k);
    bool live(void* o, Kind k);

  private:
    std::mutex mu_;
    std::unordered_set<void*> live_;
};

template <typename T> T* as(void* h) { return static_cast<T*>(static_cast<Obj*>(h)); }

void set_error(const std::string& msg);
bool cuda_ok(cudaError_t e, const char* what);

struct String : Obj {
    std::string s;
    String() : Obj(Kind::String) {}
};

struct PlatformList : Obj {
    PlatformList() : Obj(Kind::PlatformList) {}
};

struct Platform : Obj {
    String* name;
    String* vendor;
    Platform();
    ~Platform() override;
};

struct Device : Obj {
    int ordinal = 0;
    Platform* platform = nullptr;
    std::string name;
    int sms = 0;
    uint64_t mem = 0;
    int cc_major = 0, cc_minor = 0;
    Device() : Obj(Kind::Device) {}
    ~Device() override { if (platform) platform->release(); }
};

struct Context : Obj {
    int ordinal = 0;
    Device* device = nullptr;
    Context() : Obj(Kind::Context) {}
    ~Context() override { if (device) device->release(); }
};

struct Event : Obj {
    // created lazily on the device of the first stream that records it
    cudaEvent_t ev = nullptr;
    int ordinal = -1;
    std::atomic<bool> recorded{false};
    std::mutex mu;
    Event() : Obj(Kind::Event) {}
    ~Event() override;
    bool record(cudaStream_t s, int dev);
};

struct EventArr : Obj {
    std::vector<Event*> evs;
    std::mutex mu;
    EventArr() : Obj(Kind::EventArr) {}
    ~EventArr() override { for (auto* e : evs) e->release(); }
};

struct Queue : Obj {
    Context* ctx = nullptr;
    int ordinal = 0;
    cudaStream_t stream = nullptr;
    // marker counting (ClCommandQueue.cs:100-117): polled events instead of host callbacks
    std::mutex mu;
    std::deque<cudaEvent_t> pending;
    std::vector<cudaEvent_t> pool;
    int reached = 0;
    Queue() : Obj(Kind::Queue) {}
    ~Queue() override;
    void wait_list(EventArr* arr);
};

struct UserEvent : Obj {
    Context* ctx = nullptr;
    // a host-signalled gate (ClUserEvent.cs:48-121): queues wait, through a stream memory-wait,
    // on a mapped pinned flag that trigger()/the last decrement writes from the host
    volatile int* flag = nullptr;
    std::mutex mu;
    int counter = 0;
    UserEvent() : Obj(Kind::UserEvent) {}
    ~UserEvent() override;
};

struct NdRange : Obj {
    int n;
    explicit NdRange(int v) : Obj(Kind::NdRange), n(v) {}
};

struct Buffer : Obj {
    Context* ctx = nullptr;
    int ordinal = 0;
    void* dptr = nullptr;
    int64_t elems = 0;
    int elem_size = 4;
    bool zero_copy = false;  // dptr aliases mapped pinned host memory; transfers are no-ops
    bool owns = true;
    Buffer() : Obj(Kind::Buffer) {}
    ~Buffer() override;
    int64_t bytes() const { return elems * (int64_t)elem_size; }
};

struct Cubin {
    std::vector<char> image;
    std::string log;
    int err = 0;
    std::vector<std::pair<std::string, int>> kernels;  // name, number of user parameters
};

struct Program : Obj {
    Context* ctx = nullptr;
    std::shared_ptr<Cubin> cubin;
    cudaLibrary_t lib = nullptr;
    int err = 0;
    std::string log;
    Program() : Obj(Kind::Program) {}
    ~Program() override;
};

struct Kernel : Obj {
    Program* prog = nullptr;
    cudaKernel_t fn = nullptr;
    std::string name;
    int err = 0;
    int nparams = -1;            // user parameters (the hidden offset is not counted); -1 = unknown
    std::vector<Buffer*> args;   // retained
    std::mutex mu;
    Kernel() : Obj(Kind::Kernel) {}
    ~Kernel() override;
};

struct Array : Obj {
    void* raw = nullptr;
    void* head = nullptr;
    int64_t bytes = 0;
    bool pinned = false;
    Array() : Obj(Kind::Array) {}
    ~Array() override;
};

struct Timer : Obj {
    int ordinal = 0;
    cudaEvent_t a = nullptr, b = nullptr;
    bool started = false, stopped = false;
    Timer() : Obj(Kind::Timer) {}
    ~Timer() override;
};

// pinned-array table: lets createBuffer recognise zero-copy capable host pointers
bool pinned_lookup(const void* p, int64_t* bytes_out);
void pinned_add(void* p, int64_t bytes);
void pinned_remove(void* p);

// NCCL broadcast (ext_nccl.cu); nccl_broadcast returns 0 ok, <0 unavailable, >0 NCCL error
bool nccl_available();
int nccl_broadcast(const std::vector<int>& ordinals, const std::vector<void*>& ptrs, const std::vector<cudaStream_t>& streams,
                   size_t bytes, int root);
unsigned nccl_next_root(const std::vector<int>& ordinals);

// program front-end (program.cpp)
std::shared_ptr<Cubin> compile_source(const std::string& user_src, int cc_major, int cc_minor);
std::string rewrite_kernels(const std::string& src, std::vector<std::pair<std::string, int>>* kernels_out);

}  // namespace ck
