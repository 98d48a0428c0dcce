// worker.h — one device engine: context, queues, program, kernel/buffer/range caches, timing.
//
// B200-native restatement of the reference's Worker (Worker.cs:214-1735).  It talks to the GPU only through
// the KutuphaneCL C ABI (include/KutuphaneCL.h), issuing the same kinds of calls the C# Worker issues, so the
// unmodified C# layer and this C++ layer exercise the same boundary.
// Differences from the reference, all host-side:
//   * per-array flag strings (Appendix B of SURVEY.md) are parsed once into a bitmask, not re-parsed with
//     string.Contains on every transfer (Worker.cs:827-834);
//   * caches are keyed by pointers/ints, not by concatenated strings (Worker.cs:999);
//   * the load-balancer clock is a pair of CUDA events on the device's main queue (quirk Q10), with the
//     reference's host stopwatch kept for the pipelined paths, where work spans several queues.
#pragma once
#include <KutuphaneCL.h>

#include <chrono>
#include <map>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

namespace cekirdekler {

enum : uint32_t {
    F_PARTIAL = 1u << 0,  // " partial " — upload only the device's slice
    F_READ = 1u << 1,     // " read "    — (without partial) upload the whole array to every device
    F_WRITE = 1u << 2,    // " write "   — download the device's slice
    F_ALL = 1u << 3,      // " all "     — with write: whole array, from device (arrayIndex % D) only
    F_RO = 1u << 4,       // " ro "
    F_WO = 1u << 5,       // " wo "
    F_ZC = 1u << 6,       // " zc "      — zero-copy allowed if the device is in "stream" mode
};
uint32_t parseFlags(const char* readWrite);  // substring semantics of Worker.cs:827-834 / :1349-1356

// What a kernel parameter is, as far as a Worker cares: a host pointer, its element type and length.
struct ArrayRef {
    void* host = nullptr;    // base address of the host array (FastArr head or caller memory)
    int64_t length = 0;      // elements
    int clType = 0;          // ClBuffer.SizeOf code (ClBuffer.cs:266-275)
    bool fastArr = false;    // allocated by createArray (pinned + mapped): zero-copy capable
    uint32_t flags = 0;      // F_* bitmask
    int elemsPerWorkItem = 1;
    const void* identity = nullptr;  // cache key: the user-visible array object (Worker.cs:583 keys by object)
};

class Worker {
  public:
    Worker(int deviceOrdinal, const std::string& kernelSource, const std::vector<std::string>& kernelNames,
           bool noPipelining, int computeQueueConcurrency, bool streamingEnabled);
    ~Worker();
    Worker(const Worker&) = delete;
    Worker& operator=(const Worker&) = delete;

    int errorCode() const { return err_; }
    const std::string& errorMessage() const { return errMsg_; }
    const std::string& deviceName() const { return deviceName_; }
    int deviceOrdinal() const { return ordinal_; }
    int computeUnits() const { return computeUnits_; }
    uint64_t memorySize() const { return memSize_; }
    bool gddr() const { return !streaming_; }  // ClDevice.isGddr(): false when GPU_STREAM was requested

    // ---- queues (Worker.cs:80-101, :230-259) ----
    ck_handle mainQueue() const { return q_[0]; }
    ck_handle readQueue(int lane) const { return lane == 0 ? qRead_ : qRead2_; }
    ck_handle writeQueue(int lane) const { return lane == 0 ? qWrite_ : qWrite2_; }
    ck_handle computeQueue(int index1) const { return q_[index1 - 1]; }  // 1-based like commandQueue..commandQueue16
    int numComputeQueues() const { return (int)q_.size(); }
    ck_handle nextComputeQueue(int cursor);  // Worker.cs:435-437
    void finishUsedComputeQueues();          // Worker.cs:364-420
    void markQueueUsed(ck_handle q);
    ck_handle lastUsedQueue() const { return lastUsed_; }
    int countMarkersAdded() const { return markersAdded_; }
    int countMarkersReached();
    void addMarker(ck_handle q);
    void finishAll();

    // ---- caches ----
    ck_handle buffer(const ArrayRef& a);                               // Worker.cs:576-720
    ck_handle kernelWithId(const std::string& name, int computeId);    // Worker.cs:291-316
    ck_handle range(int n);                                            // Worker.cs:728-750
    void dropBuffer(const void* identity);

    // ---- hot path ----
    void kernelArgument(const std::string& kernel, const std::vector<ArrayRef>& arrs, int computeId);  // Worker.cs:990-1021
    void writeToBuffer(const std::vector<ArrayRef>& arrs, int reference, int range, bool enqueueMode, ck_handle q = nullptr);     // :821-840
    void writeToBufferWithoutPartial(const std::vector<ArrayRef>& arrs);                                                         // :847-860
    void readFromBuffer(const std::vector<ArrayRef>& arrs, int reference, int range, int deviceIndex, int numDevices,
                        bool enqueueMode, ck_handle q = nullptr);                                                                // :1343-1361
    void readFromBufferAllData(const std::vector<ArrayRef>& arrs, int deviceIndex, int numDevices);                              // :871-885 (Q2 fixed)
    int writePartialEvent(const std::vector<ArrayRef>& arrs, int reference, int range, ck_handle waitArr, ck_handle evt,
                          ck_handle q);                                                                                          // :1411-1483
    int readPartialEvent(const std::vector<ArrayRef>& arrs, int reference, int range, ck_handle waitArr, ck_handle evt,
                         ck_handle q);                                                                                           // :1500-1567
    int compute(const std::string& kernel, int reference, int globalRange, int localRange, int computeId, ck_handle q = nullptr);  // :1033-1039
    int computeRepeated(const std::string& kernel, int reference, int globalRange, int localRange, int computeId, int repeats,
                        ck_handle q = nullptr);                                                                                  // :1052-1058
    int computeRepeatedWithSyncKernel(const std::string& kernel, int reference, int globalRange, int localRange, int computeId,
                                      int repeats, const std::string& syncKernel, ck_handle q = nullptr);                        // :1072-1081
    int computeEvent(const std::string& kernel, int reference, int globalRange, int localRange, int computeId, ck_handle waitArr,
                     ck_handle evt, ck_handle q);                                                                                // :1108-1114

    // ---- balancer clock (Worker.cs:753-807) ----
    void benchmark0(int computeId) { bench_[computeId] = 0.0; }
    void startBench(bool deviceClock);
    void stopBench();
    void endBench(int computeId);
    double benchmark(int computeId) const;

    int launchErrors() const { return launchErrors_; }
    const std::string& lastLaunchError() const { return lastLaunchError_; }

    std::mutex mu;  // the reference serialises per-worker work with lock(workers[i]) (Cores.cs:751)

  private:
    void fail(int code, const std::string& msg);
    int noteLaunch(int err, const std::string& kernel);
    int launchErrors_ = 0;
    std::string lastLaunchError_;

    int ordinal_ = 0;
    bool streaming_ = false;
    int err_ = 0;
    std::string errMsg_;
    std::string deviceName_;
    int computeUnits_ = 0;
    uint64_t memSize_ = 0;
    int concurrency_ = 16;

    ck_handle platform_ = nullptr, device_ = nullptr, context_ = nullptr, program_ = nullptr;
    std::vector<ck_handle> q_;  // commandQueue, commandQueue2..16
    ck_handle qRead_ = nullptr, qWrite_ = nullptr, qRead2_ = nullptr, qWrite2_ = nullptr;
    std::vector<int> used_;     // per compute queue use counts (numComputeQueueUsed)
    ck_handle lastUsed_ = nullptr;
    int markersAdded_ = 0;

    std::map<std::pair<std::string, int>, ck_handle> kernels_;
    std::unordered_map<const void*, ck_handle> buffers_;
    std::unordered_map<int, ck_handle> ranges_;
    struct Binding { const void* identity; int elems; };
    std::map<std::pair<ck_handle, int>, Binding> bindings_;  // (kernel object, arg index) → what is bound

    std::map<int, double> bench_;
    ck_handle timer_ = nullptr;
    bool timerRunning_ = false;
    bool deviceClock_ = false;
    bool stopRecorded_ = false;
    std::chrono::steady_clock::time_point t0_;
};

}  // namespace cekirdekler
