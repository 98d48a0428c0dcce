// cores.cpp — see cores.h.  Reference: Cores.cs:60-127 (mode switches), :290-344 (construction),
// :471-963 (compute), :994-1130 (report, benchmarks, history), :1197-1980 (computePipelined).
#include "cores.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace cekirdekler {

// ------------------------------------------------------------------ DeviceThreads
DeviceThreads::DeviceThreads(int n) : n_(n) {
    for (int i = 1; i < n_; i++) threads_.emplace_back([this, i] { loop(i); });
}
DeviceThreads::~DeviceThreads() {
    {
        std::lock_guard<std::mutex> lk(mu_);
        stop_ = true;
    }
    cvStart_.notify_all();
    for (auto& t : threads_) t.join();
}
void DeviceThreads::loop(int i) {
    uint64_t seen = 0;
    for (;;) {
        const std::function<void(int)>* fn;
        {
            std::unique_lock<std::mutex> lk(mu_);
            cvStart_.wait(lk, [&] { return stop_ || epoch_ != seen; });
            if (stop_) return;
            seen = epoch_;
            fn = fn_;
        }
        (*fn)(i);
        {
            std::lock_guard<std::mutex> lk(mu_);
            if (--pending_ == 0) cvDone_.notify_all();
        }
    }
}
void DeviceThreads::run(const std::function<void(int)>& fn) {
    if (n_ <= 1) {
        if (n_ == 1) fn(0);
        return;
    }
    {
        std::lock_guard<std::mutex> lk(mu_);
        fn_ = &fn;
        pending_ = n_ - 1;
        epoch_++;
    }
    cvStart_.notify_all();
    fn(0);
    std::unique_lock<std::mutex> lk(mu_);
    cvDone_.wait(lk, [&] { return pending_ == 0; });
}

// ------------------------------------------------------------------ construction
Cores::Cores(const std::vector<int>& deviceOrdinals, const std::string& kernelSource, const std::vector<std::string>& kernelNames,
             int computeQueueConcurrency, bool noPipelining, bool streamingEnabled) {
    if (deviceOrdinals.empty()) {
        ctorError_ = "Cores: no devices selected";
        return;
    }
    workers_.resize(deviceOrdinals.size());
    threads_.reset(new DeviceThreads((int)deviceOrdinals.size()));
    // one worker per device, built concurrently (each compiles/loads the program for its own context)
    threads_->run([&](int i) {
        workers_[i].reset(new Worker(deviceOrdinals[i], kernelSource, kernelNames, noPipelining, computeQueueConcurrency, streamingEnabled));
    });
}

Cores::~Cores() {
    for (auto& w : workers_)
        if (w) w->finishAll();
    workers_.clear();
}

int Cores::errorCode() const {
    if (!ctorError_.empty()) return 1;
    int e = 0;
    for (auto& w : workers_) e += w ? w->errorCode() : 1;  // codes are summed (Cores.cs:337)
    return e;
}
std::string Cores::errorMessage() const {
    std::string m = ctorError_;
    for (auto& w : workers_)
        if (w) m += w->errorMessage();
    return m;
}
std::vector<std::string> Cores::deviceNames() const {
    std::vector<std::string> n;
    for (auto& w : workers_) n.push_back(w->deviceName());
    return n;
}

void Cores::forEachWorker(const std::function<void(int)>& fn, bool parallel) {
    if (parallel && workers_.size() > 1) threads_->run(fn);
    else
        for (int i = 0; i < (int)workers_.size(); i++) fn(i);
}

// ------------------------------------------------------------------ mode switches (Cores.cs:66-127)
void Cores::setEnqueueModeAsyncEnable(bool v) {
    enqueueModeAsync_ = v;
    if (v) asyncQueueIndex_++;
}
void Cores::setEnqueueMode(bool v) {
    if (v && !enqueueMode_)
        for (auto& w : workers_) w->startBench(false);
    const bool was = enqueueMode_;
    enqueueMode_ = v;
    if (!v) {
        forEachWorker(
            [&](int i) {
                workers_[i]->finishUsedComputeQueues();
                if (was) workers_[i]->endBench(lastUsedComputeId);
            },
            true);
    } else {
        asyncQueueIndex_ = 0;
    }
}
int Cores::countMarkers() {
    int r = 0;
    for (auto& w : workers_) r += w->countMarkersAdded();
    return r;
}
int Cores::countMarkerCallbacks() {
    int r = 0;
    for (auto& w : workers_) r += w->countMarkersReached();
    return r;
}
void Cores::flushLastUsedCommandQueue() {
    for (auto& w : workers_) flush(w->lastUsedQueue() ? w->lastUsedQueue() : w->mainQueue());
}

// ------------------------------------------------------------------ state accessors
const ComputeIdState* Cores::state(int computeId) const {
    auto it = states_.find(computeId);
    return it == states_.end() ? nullptr : &it->second;
}
std::vector<double> Cores::benchmarks(int computeId) {
    std::vector<double> b;
    for (auto& w : workers_) b.push_back(w->benchmark(computeId));
    return b;
}
void Cores::resetBalancer(int computeId) { states_.erase(computeId); }

std::vector<double> Cores::normalizedComputePowers() const {
    auto it = states_.find(lastUsedComputeId);
    if (it == states_.end()) return std::vector<double>(workers_.size(), 0.0);
    const auto& t = it->second.balance.lastThroughputs;
    std::vector<double> r(t.size(), 0.0);
    double s = 0;
    for (double v : t) s += v;
    for (size_t i = 0; i < r.size() && i < t.size(); i++) r[i] = s > 0 ? t[i] / s : 0.0;
    return r;
}
std::vector<double> Cores::normalizedGlobalRanges(int computeId) const {
    auto it = states_.find(computeId);
    if (it == states_.end()) return std::vector<double>(workers_.size(), 0.0);
    std::vector<double> r(it->second.ranges.size(), 0.0);
    double s = 0;
    for (int v : it->second.ranges) s += v;
    for (size_t i = 0; i < r.size(); i++) r[i] = s > 0 ? it->second.ranges[i] / s : 0.0;
    return r;
}

std::string Cores::performanceReport(int computeId) {
    // layout of Cores.cs:994-1063
    if (computeId == 0) {
        if (states_.empty()) return "Needs one more compute to profile. Load balancer needs multiple iterations to be useful.";
        computeId = states_.begin()->first;
    }
    auto it = states_.find(computeId);
    if (it == states_.end()) return "Error: Global range array is not ready.";
    const auto& st = it->second;
    std::string percent = "----- Load Distributions: ";
    double total = 0;
    for (int v : st.ranges) total += v;
    char buf[256];
    for (int v : st.ranges) {
        snprintf(buf, sizeof buf, " [%.1f%%] -", total > 0 ? 100.0 * v / total : 0.0);
        percent += buf;
    }
    if (percent.size() < 50) percent += std::string(50 - percent.size(), '-');
    std::string sb = "\n\nCompute-ID: " + std::to_string(computeId) + "  " + percent + "------------------------------------------------\n";
    for (int i = 0; i < (int)workers_.size(); i++) {
        std::string line = "Device " + std::to_string(i) + "(" + (workers_[i]->gddr() ? "gddr" : "stream") + "): " + workers_[i]->deviceName();
        if (line.size() > 50) line.resize(50);
        line += std::string(50 - line.size(), ' ');
        snprintf(buf, sizeof buf, " ||| time: %.2fms, workitems: %d\n", workers_[i]->benchmark(computeId),
                 st.ranges[(spmdWorld_ > 0 ? spmdRank_ : 0) + i]);
        sb += line + buf;
    }
    sb += "-----------------------------------------------------------------------------------------------------------------\n";
    return sb;
}

// ------------------------------------------------------------------ compute (Cores.cs:471-963)
void Cores::compute(const std::vector<std::string>& kernelNames, int numRepeats, const std::string& syncKernelName,
                    std::vector<ArrayRef> arrs, int globalRange, int computeId, int globalOffset, bool pipelineEnabled,
                    int numberOfPipelineStages, bool pipelineType, int localRange) {
    if (errorCode() != 0) {
        fprintf(stderr, "Cores class compute error.\n%s\n", errorMessage().c_str());
        return;
    }
    const int W = (int)workers_.size();                  // devices driven by this process
    const int D = spmdWorld_ > 0 ? spmdWorld_ : W;       // devices the range is split over
    const int base = spmdWorld_ > 0 ? spmdRank_ : 0;     // global index of local worker 0
    localRange_ = localRange;
    if (numberOfPipelineStages < 1) numberOfPipelineStages = 1;
    const bool hasSync = !syncKernelName.empty() && numRepeats > 1;

    // ---- split + balance (Cores.cs:544-613) ----
    ComputeIdState& st = states_[computeId];
    bool first = st.ranges.empty();
    if (!first) {
        first = true;
        for (int v : st.ranges) first = first && (v == 0);
    }
    if (st.ranges.empty()) {
        st.ranges.assign(D, 0);
        st.references.assign(D, 0);
        st.balance.reset(D);
    }
    std::vector<double> bench(D, 10.0);
    if (first) {
        initialSplit(globalRange, D, st.ranges);
    } else {
        for (int i = 0; i < W; i++) bench[base + i] = workers_[i]->benchmark(computeId);
        if (spmdWorld_ > 0 && exchangeBenchmarks) exchangeBenchmarks(bench);  // all-gather of per-rank device times
    }
    int step = balanceStep(st.ranges, pipelineEnabled, numberOfPipelineStages, localRange, globalRange);
    if (globalRange % step != 0) step = localRange;  // guard: the fix-up loops need total ≡ 0 (mod step)
    if (localRange > 0 && globalRange % localRange == 0)
        loadBalance(bench, smoothLoadBalancer, st.balance, globalRange, st.ranges, step, balancerMode);
    st.lastBench = bench;
    st.calls++;

    int acc = 0;
    for (int g = 0; g < D; g++) {
        st.references[g] = acc + globalOffset;
        acc += st.ranges[g];
    }
    for (auto& w : workers_) w->benchmark0(computeId);
    // from here on `ranges`/`refs` are indexed by LOCAL worker index
    const std::vector<int> ranges(st.ranges.begin() + base, st.ranges.begin() + base + W);
    const std::vector<int> refs(st.references.begin() + base, st.references.begin() + base + W);

    // ---- pipelining conditions (Cores.cs:620-652) ----
    const int blobs = numberOfPipelineStages;
    bool case1 = true;
    for (int g = 0; g < D; g++) {
        case1 = case1 && (((st.ranges[g] / blobs) % localRange) == 0);
        case1 = case1 && (st.ranges[g] >= localRange * blobs);
    }
    const bool pipelined = (numRepeats <= 1) && pipelineEnabled && case1 && !enqueueMode_ && !noComputeMode;

    if (pipelined) {
        // no inter-device synchronisation: every device pipelines its own share (Cores.cs:666-735)
        forEachWorker(
            [&](int i) {
                if (kernelNames.size() == 1) {
                    computePipelined(i, syncKernelName, kernelNames[0], arrs, numRepeats, ranges, blobs, refs, pipelineType, computeId, -1);
                } else {
                    // first kernel streams the inputs in, last kernel streams the results out; kernels in between run
                    // on the whole share.  (The reference uses kernelNames[1] for the last stage and identical
                    // ONLY_READ/ONLY_WRITE constants — quirk Q1; this uses distinct modes and the last name.)
                    computePipelined(i, syncKernelName, kernelNames.front(), arrs, numRepeats, ranges, blobs, refs, pipelineType, computeId, 0);
                    const int reps = numRepeats > 0 ? numRepeats : 1;
                    for (int r = 0; r < reps; r++) {
                        for (size_t s = 1; s + 1 < kernelNames.size(); s++)
                            workers_[i]->compute(kernelNames[s], refs[i], ranges[i], localRange, computeId);
                        if (hasSync) workers_[i]->compute(syncKernelName, 0, localRange, localRange, computeId);
                    }
                    computePipelined(i, syncKernelName, kernelNames.back(), arrs, numRepeats, ranges, blobs, refs, pipelineType, computeId, 1);
                }
            },
            true);
        lastUsedComputeId = computeId;
        return;
    }

    // ---- plain path: upload → kernels → download ----
    const bool single = (W == 1);
    const bool enq = enqueueMode_ && single;  // enqueue mode only applies on the single-device branch (Appendix G)
    ck_handle cq = nullptr;
    if (enq && enqueueModeAsync_) cq = workers_[0]->nextComputeQueue(asyncQueueIndex_);

    // An array that every device uploads whole AND some device downloads into needs the reference's barrier between the
    // upload and download phases (Cores.cs:770, "now all devices have up-to-date data"); otherwise devices run free.
    bool hostHazard = false, allPinned = true, anyWholeRead = false;
    for (const auto& a : arrs) {
        const bool wholeRead = (a.flags & F_READ) && !(a.flags & F_PARTIAL);
        if (wholeRead && (a.flags & F_WRITE)) hostHazard = true;
        if (wholeRead) anyWholeRead = true;
        if ((a.flags & (F_READ | F_PARTIAL | F_WRITE)) && !a.fastArr) allPinned = false;
    }

    auto bindArgs = [&](int i) {
        if (noComputeMode) return;
        for (const auto& k : kernelNames) workers_[i]->kernelArgument(k, arrs, computeId);
        if (hasSync) workers_[i]->kernelArgument(syncKernelName, arrs, computeId);
    };
    auto launchKernels = [&](int i) {
        if (noComputeMode) return;
        Worker& w = *workers_[i];
        if (numRepeats > 0) {
            if (kernelNames.size() > 1) {
                for (int r = 0; r < numRepeats; r++) {
                    for (const auto& k : kernelNames) w.compute(k, refs[i], ranges[i], localRange, computeId, cq);
                    if (hasSync) w.compute(syncKernelName, 0, localRange, localRange, computeId, cq);
                }
            } else if (!hasSync) {
                w.computeRepeated(kernelNames[0], refs[i], ranges[i], localRange, computeId, numRepeats, cq);
            } else {
                w.computeRepeatedWithSyncKernel(kernelNames[0], refs[i], ranges[i], localRange, computeId, numRepeats, syncKernelName, cq);
            }
        } else {
            for (const auto& k : kernelNames) w.compute(k, refs[i], ranges[i], localRange, computeId, cq);
        }
    };

    if (single) {
        // One process per GPU: a whole-array `read` input crosses ONE rank's host link (the root rotates) and reaches the
        // other ranks by ncclBroadcast over NVLink, in place of `world` host uploads.  Every rank takes part, also one
        // whose share is empty; the ranks agree on the root because they issue the same sequence of calls.
        // It pays when `world` uploads would queue up on a shared host link for longer than the collective costs: a
        // broadcast makes every rank wait for the slowest one each call, on top of NCCL's own latency.  Measured (round 1,
        // 2 MiB of positions per call): on 2 B200s the broadcast loses 40 us per call to two concurrent uploads, on 8 it
        // loses 160 us (n-body step 1.157 ms against 0.995 ms) although the box's host link is shared (107 GB/s aggregate).
        // With those numbers the break-even is near 4 MB at 8 ranks; the defaults are world >= 4 and arrays >= 16 MiB
        // (CK_SPMD_BROADCAST_MIN_WORLD / CK_SPMD_BROADCAST_MIN_BYTES override).  Smaller arrays are uploaded by every rank.
        static const long long minBytes = getenv("CK_SPMD_BROADCAST_MIN_BYTES") ? atoll(getenv("CK_SPMD_BROADCAST_MIN_BYTES")) : (16ll << 20);
        static const int minWorld = getenv("CK_SPMD_BROADCAST_MIN_WORLD") ? atoi(getenv("CK_SPMD_BROADCAST_MIN_WORLD")) : 4;
        std::vector<ArrayRef> upArrs = arrs;
        if (spmdWorld_ > 1 && spmdWorld_ >= minWorld && anyWholeRead && broadcastWholeReads && !noComputeMode &&
            ckSpmdWorld() == spmdWorld_) {
            Worker& w = *workers_[0];
            std::lock_guard<std::mutex> lk(w.mu);
            ck_handle q = cq ? cq : w.mainQueue();
            bool timed = false;
            for (size_t a = 0; a < arrs.size(); a++) {
                const auto& ar = arrs[a];
                if (!((ar.flags & F_READ) && !(ar.flags & F_PARTIAL)) || (ar.flags & F_WRITE)) continue;
                ck_handle b = w.buffer(ar);
                if (!b || ckBufferIsZeroCopy(b) || ckBufferBytes(b) < minBytes) continue;
                if (!timed && !enq && ranges[0] > 0) {
                    bindArgs(0);
                    w.startBench(deviceClock);
                    timed = true;
                }
                const int root = (int)(spmdBroadcastCalls_++ % (unsigned)spmdWorld_);
                if (root == spmdRank_) ::writeToBuffer(q, b, ar.host);
                if (ckSpmdBroadcast(q, b, root) == 0) upArrs[a].flags &= ~(uint32_t)(F_READ | F_PARTIAL);
            }
            if (ranges[0] > 0) {
                if (!timed) {
                    bindArgs(0);
                    if (!enq) w.startBench(deviceClock);
                }
                w.writeToBuffer(upArrs, refs[0], ranges[0], /*enqueueMode=*/true, cq);
                launchKernels(0);
                w.readFromBuffer(arrs, refs[0], ranges[0], base, D, /*enqueueMode=*/true, cq);
                if (!enq) {
                    w.stopBench();
                    finish(w.mainQueue());
                    w.endBench(computeId);
                } else {
                    w.markQueueUsed(w.mainQueue());
                    if (fineGrainedQueueControl && enqueueModeAsync_) w.addMarker(w.lastUsedQueue() ? w.lastUsedQueue() : w.mainQueue());
                }
            } else if (!enq) {
                finish(q);
            }
        } else if (ranges[0] > 0) {
            Worker& w = *workers_[0];
            std::lock_guard<std::mutex> lk(w.mu);
            bindArgs(0);
            if (!enq) w.startBench(deviceClock);
            w.writeToBuffer(arrs, refs[0], ranges[0], /*enqueueMode=*/true, cq);  // in-order queue: no host sync needed before the kernel
            launchKernels(0);
            w.readFromBuffer(arrs, refs[0], ranges[0], base, D, /*enqueueMode=*/true, cq);
            if (!enq) {
                w.stopBench();  // the span ends when the download does, not when the host thread wakes up from the wait
                finish(w.mainQueue());
                w.endBench(computeId);
            }
            if (enq) {
                w.markQueueUsed(w.mainQueue());
                if (fineGrainedQueueControl && enqueueModeAsync_) w.addMarker(w.lastUsedQueue() ? w.lastUsedQueue() : w.mainQueue());
            }
        }
    } else {
        // whole-array `read` inputs: one host-link upload + NVLink fan-out instead of D host uploads
        std::vector<bool> broadcasted(arrs.size(), false);
        if (anyWholeRead && broadcastWholeReads) {
            std::vector<int> active;
            for (int i = 0; i < W; i++)
                if (ranges[i] > 0) active.push_back(i);
            if (active.size() > 1) {
                for (size_t a = 0; a < arrs.size(); a++) {
                    const auto& ar = arrs[a];
                    if (!((ar.flags & F_READ) && !(ar.flags & F_PARTIAL))) continue;
                    std::vector<ck_handle> qs, bs;
                    bool ok = true;
                    for (int i : active) {
                        ck_handle b = workers_[i]->buffer(ar);
                        if (!b || ckBufferIsZeroCopy(b)) ok = false;
                        qs.push_back(workers_[i]->mainQueue());
                        bs.push_back(b);
                    }
                    if (ok && ckBroadcastUpload(qs.data(), bs.data(), (int)qs.size(), ar.host) == 0) broadcasted[a] = true;
                }
            }
        }
        std::vector<ArrayRef> upArrs = arrs;  // arrays already broadcast are not uploaded again per device
        for (size_t a = 0; a < arrs.size(); a++)
            if (broadcasted[a]) upArrs[a].flags &= ~(uint32_t)(F_READ | F_PARTIAL);

        auto phase1 = [&](int i) {
            if (ranges[i] <= 0) return;
            Worker& w = *workers_[i];
            std::lock_guard<std::mutex> lk(w.mu);
            bindArgs(i);
            w.startBench(deviceClock);
            w.writeToBuffer(upArrs, refs[i], ranges[i], /*enqueueMode=*/!hostHazard);
        };
        auto phase23 = [&](int i) {
            if (ranges[i] <= 0) return;
            Worker& w = *workers_[i];
            std::lock_guard<std::mutex> lk(w.mu);
            launchKernels(i);
            w.readFromBuffer(arrs, refs[i], ranges[i], base + i, D, /*enqueueMode=*/true);
        };
        auto drain = [&](int i) {
            if (ranges[i] <= 0) return;
            workers_[i]->endBench(computeId);  // device clock: waits for the stop event, i.e. for the download
            finish(workers_[i]->mainQueue());
        };
        // pageable host memory makes copies block, and the host stopwatch must not see other devices' waits:
        // in both cases give each device its own thread (as the reference's Parallel.For does)
        const bool threaded = !allPinned || !deviceClock;
        if (hostHazard) {
            forEachWorker(phase1, threaded);
            forEachWorker(phase23, threaded);
        } else {
            forEachWorker([&](int i) { phase1(i); phase23(i); }, threaded);
        }
        if (!threaded)
            for (int i = 0; i < W; i++)
                if (ranges[i] > 0) workers_[i]->stopBench();  // every device's stop event goes in before the first wait
        forEachWorker(drain, threaded);
    }
    lastUsedComputeId = computeId;
}

// ------------------------------------------------------------------ computePipelined (Cores.cs:1197-1980)
// readWrite: -1 = upload+compute+download, 0 = upload+compute only, 1 = compute+download only.
void Cores::computePipelined(int i, const std::string& syncKernelName, const std::string& kernelName, const std::vector<ArrayRef>& arrs,
                             int numRepeats, const std::vector<int>& ranges, int blobs, const std::vector<int>& offsets, bool pipelineType,
                             int computeId, int readWrite) {
    if (ranges[i] <= 0) return;
    Worker& w = *workers_[i];
    std::lock_guard<std::mutex> lk(w.mu);
    const bool doUp = (readWrite == -1 || readWrite == 0);
    const bool doDown = (readWrite == -1 || readWrite == 1);
    const int D = spmdWorld_ > 0 ? spmdWorld_ : (int)workers_.size();
    const int gi = (spmdWorld_ > 0 ? spmdRank_ : 0) + i;
    const int chunk = ranges[i] / blobs;
    const int off = offsets[i];

    w.kernelArgument(kernelName, arrs, computeId);
    if (!syncKernelName.empty() && numRepeats > 1) w.kernelArgument(syncKernelName, arrs, computeId);
    if (doUp) w.startBench(false);  // work spans several queues: host stopwatch, as in the reference
    if (doUp) w.writeToBufferWithoutPartial(arrs);

    bool downloaded = false;
    if (pipelineType == PIPELINE_EVENT) {
        // Two lanes, each a 3-queue software pipeline advanced in lock-step.
        const int half = blobs / 2;
        const int steps = half + 3;
        struct Lane {
            std::vector<ck_handle> eRead, eCompute, eWrite;
        } lane[2];
        for (auto& L : lane)
            for (int j = 0; j < steps; j++) {
                L.eRead.push_back(createEventArr(0));
                L.eCompute.push_back(createEventArr(0));
                L.eWrite.push_back(createEventArr(0));
            }
        for (int j = 0; j < half + 2; j++) {
            if (j <= half - 1 && doUp) {
                for (int l = 0; l < 2; l++) {
                    ck_handle e = createEvent();
                    int did = w.writePartialEvent(arrs, off + chunk * j + l * (ranges[i] / 2), chunk, lane[l].eRead[j], e, w.readQueue(l));
                    if (did) {
                        addToEventArr(lane[l].eCompute[j + 1], e, 0);
                        addToEventArr(lane[l].eWrite[j + 1], e, 1);
                    }
                    deleteEvent(e);
                }
            }
            if (j >= 1 && j <= half) {
                for (int l = 0; l < 2; l++) {
                    ck_handle e = createEvent();
                    w.computeEvent(kernelName, off + chunk * (j - 1) + l * (ranges[i] / 2), chunk, localRange_, computeId, lane[l].eCompute[j], e,
                                   l == 0 ? w.mainQueue() : (w.numComputeQueues() > 1 ? w.computeQueue(2) : w.mainQueue()));
                    addToEventArr(lane[l].eRead[j + 1], e, 0);
                    addToEventArr(lane[l].eWrite[j + 1], e, 1);
                    deleteEvent(e);
                }
            }
            if (j >= 2 && doDown) {
                for (int l = 0; l < 2; l++) {
                    ck_handle e = createEvent();
                    int did = w.readPartialEvent(arrs, off + chunk * (j - 2) + l * (ranges[i] / 2), chunk, lane[l].eWrite[j], e, w.writeQueue(l));
                    if (did) {
                        downloaded = true;
                        addToEventArr(lane[l].eRead[j + 1], e, 0);
                        addToEventArr(lane[l].eCompute[j + 1], e, 1);
                    }
                    deleteEvent(e);
                }
            }
        }
        if (doDown) {
            ck_handle cq2 = w.numComputeQueues() > 1 ? w.computeQueue(2) : w.mainQueue();
            for (ck_handle q : {w.writeQueue(0), w.mainQueue(), w.readQueue(0), w.writeQueue(1), cq2, w.readQueue(1)}) flush(q);
            if (downloaded) {
                finish(w.writeQueue(0));
                finish(w.writeQueue(1));
            }
            // both lanes' kernels are done before a `write all` array is fetched whole on the main queue (the slice
            // downloads above already waited for them, so these two waits cost nothing in the usual case)
            finish(w.mainQueue());
            finish(cq2);
            w.readFromBufferAllData(arrs, gi, D);
        }
        for (auto& L : lane)
            for (int j = 0; j < steps; j++) {
                deleteEventArr(L.eRead[j]);
                deleteEventArr(L.eCompute[j]);
                deleteEventArr(L.eWrite[j]);
            }
    } else {
        // Driver-controlled: chunk k → in-order queue k % 16, upload/kernel/download back to back, no events
        // (Cores.cs:1368-1858); overlap between queues is left to the hardware copy engines and scheduler.
        const int nq = w.numComputeQueues();
        for (int k = 0; k < blobs; k++) {
            ck_handle q = w.computeQueue((k % std::min(16, nq)) + 1);
            if (doUp) w.writePartialEvent(arrs, off + chunk * k, chunk, nullptr, nullptr, q);
            w.computeEvent(kernelName, off + chunk * k, chunk, localRange_, computeId, nullptr, nullptr, q);
            if (doDown) w.readPartialEvent(arrs, off + chunk * k, chunk, nullptr, nullptr, q);
        }
        if (doDown) {
            for (int k = 1; k <= std::min(16, nq); k++) flush(w.computeQueue(k));
            for (int k = 1; k <= std::min(16, nq); k++) finish(w.computeQueue(k));
            w.readFromBufferAllData(arrs, gi, D);   // `write all`: whole array, after every chunk's kernel has finished
        }
    }
    if (!doDown) w.finishAll();  // upload-only stage of a multi-kernel pipeline: later kernels run on the main queue
    if (doDown) w.endBench(computeId);
}

}  // namespace cekirdekler
