// worker.cpp — see worker.h.  Reference: Worker.cs (line ranges cited per method in the header).
#include "worker.h"

#include <cstdio>
#include <cstring>

namespace cekirdekler {

uint32_t parseFlags(const char* s) {
    // The reference tests tokens by substring (Worker.cs:827-834): "partial", "read", "write", "all", "ro", "wo", "zc".
    // A faithful Contains("ro")/("wo") would also fire on unrelated text, so the short tokens are matched as the
    // space-padded words ClArray emits (ClArray.cs:1685-1703) or as whole words in hand-written strings.
    uint32_t f = 0;
    if (!s) return 0;
    std::string t(s);
    auto has = [&](const char* w) { return t.find(w) != std::string::npos; };
    auto word = [&](const char* w) {
        std::string padded = " " + t + " ";
        return padded.find(std::string(" ") + w + " ") != std::string::npos;
    };
    if (has("partial")) f |= F_PARTIAL;
    if (has("read")) f |= F_READ;
    if (has("write")) f |= F_WRITE;
    if (has("all")) f |= F_ALL;
    if (word("ro")) f |= F_RO;
    if (word("wo")) f |= F_WO;
    if (word("zc")) f |= F_ZC;
    return f;
}

static int clTypeSize(int t) {
    static const int sz[7] = {4, 8, 4, 8, 2, 1, 4};  // ClBuffer.SizeOf (ClBuffer.cs:266-275)
    return (t >= 0 && t < 7) ? sz[t] : 0;
}

void Worker::fail(int code, const std::string& msg) {
    err_ += code ? code : 1;
    errMsg_ += msg + "\n----\n";
}

static ck_handle makeString(const std::string& s) {
    ck_handle h = createString();
    writeToString(h, s.c_str());
    return h;
}

Worker::Worker(int deviceOrdinal, const std::string& kernelSource, const std::vector<std::string>& kernelNames,
               bool noPipelining, int computeQueueConcurrency, bool streamingEnabled)
    : ordinal_(deviceOrdinal), streaming_(streamingEnabled), concurrency_(computeQueueConcurrency) {
    ck_handle list = platformList();
    platform_ = createPlatform(list, 0);
    deletePlatformList(list);
    device_ = createDevice(platform_, CODE_GPU(), deviceOrdinal);
    if (!device_) {
        fail(1, std::string("Worker: cannot open CUDA device ") + std::to_string(deviceOrdinal) + ": " + ckGetLastError());
        return;
    }
    {
        ck_handle s = makeString(" ");
        getDeviceName(device_, s);
        deviceName_ = readFromString(s);
        deleteString(s);
    }
    computeUnits_ = deviceComputeUnits(device_);
    memSize_ = deviceMemSize(device_);
    context_ = createContext(platform_, device_);
    if (!context_) {
        fail(1, std::string("Worker: cannot create context: ") + ckGetLastError());
        return;
    }
    // commandQueue, then Read/Write, then 2..16 unless noPipelining, then Read2/Write2 (Worker.cs:230-259)
    q_.push_back(createCommandQueue(context_, device_, 0));
    qRead_ = createCommandQueue(context_, device_, 0);
    qWrite_ = createCommandQueue(context_, device_, 0);
    if (!noPipelining)
        for (int i = 2; i <= 16; i++) q_.push_back(createCommandQueue(context_, device_, 0));
    qRead2_ = createCommandQueue(context_, device_, 0);
    qWrite2_ = createCommandQueue(context_, device_, 0);
    used_.assign(16, 0);
    timer_ = ckCreateTimer(context_);

    ck_handle src = makeString(kernelSource);
    program_ = createProgram(context_, device_, src);
    deleteString(src);
    int perr = getProgramErr(program_);
    char* log = readProgramErrString(program_);
    if (perr) fail(perr, std::string(log ? log : ""));
    free(log);
    if (perr) return;
    for (const auto& name : kernelNames) {
        ck_handle k = kernelWithId(name, -0x7fffffff);  // the "compile test" object (Worker.cs:274-279)
        int kerr = k ? getKernelErr(k) : 1;
        if (kerr) fail(kerr, "kernel not found in program: \"" + name + "\"");
    }
}

Worker::~Worker() {
    finishAll();
    for (auto& kv : kernels_) deleteKernel(kv.second);
    for (auto& kv : buffers_) deleteBuffer(kv.second);
    for (auto& kv : ranges_) deleteNdRange(kv.second);
    if (timer_) ckDeleteTimer(timer_);
    if (program_) deleteProgram(program_);
    for (auto q : q_) deleteCommandQueue(q);
    for (auto q : {qRead_, qWrite_, qRead2_, qWrite2_})
        if (q) deleteCommandQueue(q);
    if (context_) deleteContext(context_);
    if (device_) deleteDevice(device_);
    if (platform_) deletePlatform(platform_);
}

void Worker::finishAll() {
    for (auto q : q_) finish(q);
    for (auto q : {qRead_, qWrite_, qRead2_, qWrite2_})
        if (q) finish(q);
}

ck_handle Worker::nextComputeQueue(int cursor) {
    int c = (concurrency_ > 16 || concurrency_ < 1) ? 16 : concurrency_;
    int idx = cursor % c;
    if (idx < 0 || idx >= (int)q_.size()) idx = 0;  // noPipelining workers only own commandQueue
    used_[idx]++;
    lastUsed_ = q_[idx];
    return q_[idx];
}

void Worker::markQueueUsed(ck_handle q) {
    for (size_t i = 0; i < q_.size(); i++)
        if (q_[i] == q) used_[i]++;
}

void Worker::finishUsedComputeQueues() {
    std::vector<ck_handle> list;
    for (size_t i = 0; i < q_.size(); i++)
        if (used_[i] > 0) {
            used_[i] = 0;
            list.push_back(q_[i]);
        }
    waitN(list.empty() ? nullptr : list.data(), q_[0], (int)list.size());
}

void Worker::addMarker(ck_handle q) {
    markersAdded_++;
    addMarkerToCommandQueue(q);
}

int Worker::countMarkersReached() {
    int r = 0;
    for (auto q : q_) r += getMarkerCounterOfCommandQueue(q);  // queues that do not exist are skipped (quirk Q16)
    return r;
}

ck_handle Worker::buffer(const ArrayRef& a) {
    auto it = buffers_.find(a.identity);
    if (it != buffers_.end()) {
        // same key but another shape (memory re-used by a different array): rebuild
        if (ckBufferBytes(it->second) == a.length * (int64_t)clTypeSize(a.clType)) return it->second;
        dropBuffer(a.identity);
    }
    // isGddr = device.isGddr() || !zc  (Worker.cs:578); only FastArr memory can be aliased (Worker.cs:611)
    const bool gddrBuffer = gddr() || !(a.flags & F_ZC) || !a.fastArr;
    ck_handle b = createBuffer(context_, (int32_t)a.length, a.clType, a.fastArr ? 0 : 1, a.fastArr ? a.host : nullptr,
                               gddrBuffer ? 1 : 0, (a.flags & F_RO) ? 1 : 0, (a.flags & F_WO) ? 1 : 0);
    if (!b) {
        fail(1, std::string("createBuffer failed: ") + ckGetLastError());
        return nullptr;
    }
    buffers_[a.identity] = b;
    return b;
}

void Worker::dropBuffer(const void* identity) {
    auto it = buffers_.find(identity);
    if (it == buffers_.end()) return;
    for (auto b = bindings_.begin(); b != bindings_.end();)
        b = (b->second.identity == identity) ? bindings_.erase(b) : std::next(b);
    deleteBuffer(it->second);
    buffers_.erase(it);
}

ck_handle Worker::kernelWithId(const std::string& name, int computeId) {
    auto key = std::make_pair(name, computeId);
    auto it = kernels_.find(key);
    if (it != kernels_.end()) return it->second;
    ck_handle s = makeString(name);
    ck_handle k = createKernel(program_, s);
    deleteString(s);
    kernels_[key] = k;
    return k;
}

ck_handle Worker::range(int n) {
    auto it = ranges_.find(n);
    if (it != ranges_.end()) return it->second;
    ck_handle r = createNdRange(n);
    ranges_[n] = r;
    return r;
}

void Worker::kernelArgument(const std::string& kernel, const std::vector<ArrayRef>& arrs, int computeId) {
    ck_handle k = kernelWithId(kernel, computeId);
    for (int i = 0; i < (int)arrs.size(); i++) {
        auto key = std::make_pair(k, i);
        auto it = bindings_.find(key);
        if (it != bindings_.end() && it->second.identity == arrs[i].identity && it->second.elems == arrs[i].elemsPerWorkItem)
            continue;  // already bound to this array
        ck_handle b = buffer(arrs[i]);
        if (!b) continue;
        setKernelArgument(k, b, i);
        bindings_[key] = Binding{arrs[i].identity, arrs[i].elemsPerWorkItem};
    }
}

void Worker::writeToBuffer(const std::vector<ArrayRef>& arrs, int reference, int range, bool enqueueMode, ck_handle q) {
    ck_handle cq = q ? q : q_[0];
    for (const auto& a : arrs) {
        if (a.flags & F_PARTIAL) {
            writeToBufferRanged(cq, buffer(a), reference * a.elemsPerWorkItem, range * a.elemsPerWorkItem, a.host);
            continue;
        }
        if (a.flags & F_READ) ::writeToBuffer(cq, buffer(a), a.host);
    }
    if (!enqueueMode) finish(q_[0]);
}

void Worker::writeToBufferWithoutPartial(const std::vector<ArrayRef>& arrs) {
    for (const auto& a : arrs)
        if ((a.flags & F_READ) && !(a.flags & F_PARTIAL)) ::writeToBuffer(q_[0], buffer(a), a.host);
    finish(q_[0]);
}

void Worker::readFromBuffer(const std::vector<ArrayRef>& arrs, int reference, int range, int deviceIndex, int numDevices,
                            bool enqueueMode, ck_handle q) {
    ck_handle cq = q ? q : q_[0];
    for (int i = 0; i < (int)arrs.size(); i++) {
        const auto& a = arrs[i];
        if ((a.flags & F_WRITE) && !(a.flags & F_ALL))
            readFromBufferRanged(cq, buffer(a), reference * a.elemsPerWorkItem, range * a.elemsPerWorkItem, a.host);
        else if ((a.flags & F_WRITE) && (a.flags & F_ALL) && deviceIndex == (i % numDevices))
            ::readFromBuffer(cq, buffer(a), a.host);
    }
    if (!enqueueMode) finish(q_[0]);
}

void Worker::readFromBufferAllData(const std::vector<ArrayRef>& arrs, int deviceIndex, int numDevices) {
    // The reference issues a host→device write here (Worker.cs:878-881, quirk Q2); "write all" means
    // device→host, so this restatement downloads.
    for (int i = 0; i < (int)arrs.size(); i++) {
        const auto& a = arrs[i];
        if ((a.flags & F_WRITE) && (a.flags & F_ALL) && deviceIndex == (i % numDevices)) ::readFromBuffer(q_[0], buffer(a), a.host);
    }
    finish(q_[0]);
}

// With several partial arrays the first transfer carries the wait list and the last one signals the event
// (Worker.cs:1444-1476); the ones in between are ordered by the in-order queue.
int Worker::writePartialEvent(const std::vector<ArrayRef>& arrs, int reference, int range, ck_handle waitArr, ck_handle evt,
                              ck_handle q) {
    int first = -1, last = -1;
    for (int i = 0; i < (int)arrs.size(); i++)
        if (arrs[i].flags & F_PARTIAL) {
            if (first < 0) first = i;
            last = i;
        }
    if (first < 0) return 0;
    for (int i = first; i <= last; i++) {
        const auto& a = arrs[i];
        if (!(a.flags & F_PARTIAL)) continue;
        writeToBufferRangedEvent(q, buffer(a), reference * a.elemsPerWorkItem, range * a.elemsPerWorkItem, a.host,
                                 i == first ? waitArr : nullptr, i == last ? evt : nullptr);
    }
    return 1;
}

int Worker::readPartialEvent(const std::vector<ArrayRef>& arrs, int reference, int range, ck_handle waitArr, ck_handle evt,
                             ck_handle q) {
    int first = -1, last = -1;
    for (int i = 0; i < (int)arrs.size(); i++)
        if ((arrs[i].flags & F_WRITE) && !(arrs[i].flags & F_ALL)) {
            if (first < 0) first = i;
            last = i;
        }
    if (first < 0) return 0;
    for (int i = first; i <= last; i++) {
        const auto& a = arrs[i];
        if (!((a.flags & F_WRITE) && !(a.flags & F_ALL))) continue;
        readFromBufferRangedEvent(q, buffer(a), reference * a.elemsPerWorkItem, range * a.elemsPerWorkItem, a.host,
                                  i == first ? waitArr : nullptr, i == last ? evt : nullptr);
    }
    return 1;
}

int Worker::compute(const std::string& kernel, int reference, int globalRange, int localRange, int computeId, ck_handle q) {
    ck_handle k = kernelWithId(kernel, computeId);
    return noteLaunch(::compute(q ? q : q_[0], k, range(reference), range(globalRange), range(localRange)), kernel);
}

int Worker::noteLaunch(int err, const std::string& kernel) {
    if (err) {
        launchErrors_++;
        lastLaunchError_ = "kernel \"" + kernel + "\" on device " + std::to_string(ordinal_) + ": " + ckGetLastError();
        fprintf(stderr, "Cekirdekler launch error: %s\n", lastLaunchError_.c_str());
    }
    return err;
}

int Worker::computeRepeated(const std::string& kernel, int reference, int globalRange, int localRange, int computeId, int repeats,
                            ck_handle q) {
    return noteLaunch(::computeRepeated(q ? q : q_[0], kernelWithId(kernel, computeId), range(reference), range(globalRange),
                                        range(localRange), repeats), kernel);
}

int Worker::computeRepeatedWithSyncKernel(const std::string& kernel, int reference, int globalRange, int localRange, int computeId,
                                          int repeats, const std::string& syncKernel, ck_handle q) {
    return noteLaunch(::computeRepeatedWithSyncKernel(q ? q : q_[0], kernelWithId(kernel, computeId), range(reference), range(globalRange),
                                                      range(localRange), repeats, kernelWithId(syncKernel, computeId), range(0)), kernel);
}

int Worker::computeEvent(const std::string& kernel, int reference, int globalRange, int localRange, int computeId, ck_handle waitArr,
                         ck_handle evt, ck_handle q) {
    return noteLaunch(::computeEvent(q, kernelWithId(kernel, computeId), range(reference), range(globalRange), range(localRange), waitArr, evt), kernel);
}

void Worker::startBench(bool deviceClock) {
    deviceClock_ = deviceClock && timer_;
    if (deviceClock_) ckTimerStart(timer_, q_[0]);
    else t0_ = std::chrono::steady_clock::now();
    timerRunning_ = true;
}

void Worker::stopBench() {
    // Records the end of the device span now, without waiting for it.  When one host thread drives several devices it must
    // do this for ALL of them before it blocks on any: an event's time stamp cannot be earlier than the moment it was
    // enqueued, so a stop event recorded after waiting for device 0 would charge device 0's run time to device 1.
    if (timerRunning_ && deviceClock_ && !stopRecorded_) {
        ckTimerStop(timer_, q_[0]);
        stopRecorded_ = true;
    }
}

void Worker::endBench(int computeId) {
    if (!timerRunning_) return;
    timerRunning_ = false;
    double ms;
    if (deviceClock_) {
        if (!stopRecorded_) ckTimerStop(timer_, q_[0]);
        stopRecorded_ = false;
        ms = ckTimerElapsedMs(timer_);
        if (ms < 0) ms = 0;
    } else {
        ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0_).count();
    }
    bench_[computeId] += ms;
}

double Worker::benchmark(int computeId) const {
    auto it = bench_.find(computeId);
    return it == bench_.end() ? 0.0 : it->second;
}

}  // namespace cekirdekler
