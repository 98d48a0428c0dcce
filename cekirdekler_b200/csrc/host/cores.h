// cores.h — the scheduler core: range split, iterative balancing, 3-phase / pipelined execution.
//
// B200-native restatement of the reference's Cores (Cores.cs:156-1982): `compute()` splits one kernel's 1-D
// global range over the selected devices, rebalances the split per compute-id from the last call's device
// times (balance.h), then runs upload → kernels → download per device, either plainly (Cores.cs:747-949) or
// chunked into `blobs` pipelined pieces in the event-driven or driver-driven mode (Cores.cs:1197-1980).
#pragma once
#include "balance.h"
#include "worker.h"

#include <condition_variable>
#include <functional>
#include <memory>
#include <thread>

namespace cekirdekler {

// Persistent fork/join helper standing in for the reference's Parallel.For over workers (Cores.cs:671, :747).
class DeviceThreads {
  public:
    explicit DeviceThreads(int n);
    ~DeviceThreads();
    void run(const std::function<void(int)>& fn);  // fn(i) for i in [0,n) — i==0 on the caller, rest on pool threads
    int size() const { return n_; }

  private:
    void loop(int i);
    int n_;
    std::vector<std::thread> threads_;
    std::mutex mu_;
    std::condition_variable cvStart_, cvDone_;
    const std::function<void(int)>* fn_ = nullptr;
    uint64_t epoch_ = 0;
    int pending_ = 0;
    bool stop_ = false;
};

struct ComputeIdState {
    std::vector<int> ranges;      // Cores.globalRanges[computeId]
    std::vector<int> references;  // Cores.globalReferences[computeId]
    BalanceState balance;
    std::vector<double> lastBench;
    int calls = 0;
};

class Cores {
  public:
    static constexpr bool PIPELINE_EVENT = true;    // Cores.cs:416-423
    static constexpr bool PIPELINE_DRIVER = false;

    Cores(const std::vector<int>& deviceOrdinals, const std::string& kernelSource, const std::vector<std::string>& kernelNames,
          int computeQueueConcurrency, bool noPipelining, bool streamingEnabled);
    ~Cores();

    // Cores.cs:471-963.  `arrs[i].flags/elemsPerWorkItem` carry what the reference passes as readWrite[]/elementsPerWorkItem[].
    void compute(const std::vector<std::string>& kernelNames, int numRepeats, const std::string& syncKernelName,
                 std::vector<ArrayRef> arrs, int globalRange, int computeId, int globalOffset = 0, bool pipelineEnabled = false,
                 int numberOfPipelineStages = 4, bool pipelineType = PIPELINE_EVENT, int localRange = 256);

    int numberOfDevices() const { return (int)workers_.size(); }
    Worker& worker(int i) { return *workers_[i]; }
    int errorCode() const;
    std::string errorMessage() const;
    std::vector<std::string> deviceNames() const;
    std::string performanceReport(int computeId);
    std::vector<double> benchmarks(int computeId);
    const ComputeIdState* state(int computeId) const;
    std::vector<double> normalizedComputePowers() const;          // ClNumberCruncher.cs:254-270
    std::vector<double> normalizedGlobalRanges(int computeId) const;  // ClNumberCruncher.cs:277-294
    void resetBalancer(int computeId);

    // ---- mode switches mirrored from ClNumberCruncher/Cores (Cores.cs:66-127) ----
    void setEnqueueMode(bool v);
    bool enqueueMode() const { return enqueueMode_; }
    void setEnqueueModeAsyncEnable(bool v);
    bool enqueueModeAsyncEnable() const { return enqueueModeAsync_; }
    bool noComputeMode = false;
    bool fineGrainedQueueControl = false;
    bool smoothLoadBalancer = true;
    BalancerMode balancerMode = BalancerMode::Reference;
    bool broadcastWholeReads = true;  // multi-device: upload whole-`read` arrays once and fan out over NVLink
    bool deviceClock = true;  // time the balancer with CUDA events (false: host stopwatch as in the reference)
    int countMarkers();
    int countMarkerCallbacks();
    void flushLastUsedCommandQueue();
    int lastUsedComputeId = -1;

    // SPMD front-end (one process per GPU, e.g. under torchrun): this process drives ONE device that plays global
    // device index `rank` of `world`.  Every rank runs the same deterministic balancer on the all-gathered device
    // times, so all ranks agree on the split without any data-path collective.
    void setSpmd(int rank, int world) { spmdRank_ = rank; spmdWorld_ = world; }
    int spmdRank() const { return spmdRank_; }
    int spmdWorld() const { return spmdWorld_; }
    // in/out: vector of `world` device times with this rank's entry filled in; must return with all entries filled
    std::function<void(std::vector<double>&)> exchangeBenchmarks;

  private:
    void computePipelined(int i, const std::string& syncKernelName, const std::string& kernelName, const std::vector<ArrayRef>& arrs,
                          int numRepeats, const std::vector<int>& ranges, int blobs, const std::vector<int>& offsets, bool pipelineType,
                          int computeId, int readWrite);
    void forEachWorker(const std::function<void(int)>& fn, bool parallel);

    std::vector<std::unique_ptr<Worker>> workers_;
    std::unique_ptr<DeviceThreads> threads_;
    std::map<int, ComputeIdState> states_;
    int localRange_ = 256;
    bool enqueueMode_ = false;
    bool enqueueModeAsync_ = false;
    int asyncQueueIndex_ = 0;
    int spmdRank_ = 0, spmdWorld_ = 0;
    unsigned spmdBroadcastCalls_ = 0;  // rotates the root of rank-to-rank broadcasts; advances identically on every rank
    std::chrono::steady_clock::time_point enqueueT0_;
    std::string ctorError_;
};

}  // namespace cekirdekler
