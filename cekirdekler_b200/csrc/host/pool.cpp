// pool.cpp — see pool.h.  Reference: ClPipeline.cs:3331-3520 (ClTask), :3607-3783 (ClTaskPool), :3872-4700 (ClDevicePool,
// producer produceTasksComputeAtWill :4132-4312), :4700-5078 (DevicePoolThread, consumer consumeTasksComputeAtWill :4841-5047).
#include "pool.h"

#include <cstdio>

namespace cekirdekler {

int ClTask::compute(ClNumberCruncher& cruncher) const {
    ClParameterGroup g;
    for (auto* a : arrays) g.nextParam(a);
    return g.compute(cruncher, computeId, kernelNamesString, globalRange, localRange, ofsetGlobalRange, pipeline, pipelineType,
                     pipelineBlobs, readWrite.empty() ? nullptr : &readWrite, elementsPerItem.empty() ? nullptr : &elementsPerItem);
}

ClDevicePool::ClDevicePool(ClDevicePoolType poolType, const std::string& kernelCodeToCompile, bool fineGrainedQueueControl,
                           int maxQueuesPerDevice)
    : type_(poolType), kernelCode_(kernelCodeToCompile), fineGrained_(fineGrainedQueueControl) {
    maxQueues_ = std::min(16, std::max(1, maxQueuesPerDevice));
    // in-flight limit per device (ClPipeline.cs:3963-3970)
    deviceQueueLimit_ = (poolType == DEVICE_COMPUTE_AT_WILL ? 2 : 25) * maxQueues_;
    producer_ = std::thread([this] { producerLoop(); });
}

ClDevicePool::~ClDevicePool() { dispose(); }

void ClDevicePool::addDevice(int deviceOrdinal) {
    auto d = std::make_unique<Device>();
    d->ordinal = deviceOrdinal;
    d->cruncher.reset(new ClNumberCruncher({deviceOrdinal}, kernelCode_, false, maxQueues_));
    Device* raw = d.get();
    {
        std::lock_guard<std::mutex> lk(mu_);
        devices_.push_back(std::move(d));
    }
    raw->thread = std::thread([this, raw] { consumerLoop(raw); });
}

int ClDevicePool::errorCode() const {
    std::lock_guard<std::mutex> lk(mu_);
    int e = 0;
    for (auto& d : devices_) e += d->cruncher->errorCode();
    return e;
}
std::string ClDevicePool::errorMessage() const {
    std::lock_guard<std::mutex> lk(mu_);
    std::string m;
    for (auto& d : devices_) m += d->cruncher->errorMessage();
    return m;
}
std::vector<int> ClDevicePool::tasksPerDevice() const {
    std::lock_guard<std::mutex> lk(mu_);
    std::vector<int> r;
    for (auto& d : devices_) r.push_back(d->consumed);
    return r;
}

void ClDevicePool::enqueueTaskPool(const ClTaskPool& pool) {
    std::lock_guard<std::mutex> lk(mu_);
    if (!timing_) {
        t0_ = std::chrono::steady_clock::now();
        timing_ = true;
    }
    pools_.push_back(pool);  // a copy: "duplicates the pool" (ClPipeline.cs:4400-4409)
    cvPools_.notify_all();
}

int ClDevicePool::loadLocked(const Device& d) const {
    int load = (int)d.priv.size() + d.inFlight;
    if (fineGrained_) load += d.cruncher->countMarkersRemaining();
    return load;
}
int ClDevicePool::leastLoadedLocked() const {
    int best = 0, bestLoad = 1 << 30;
    for (int i = 0; i < (int)devices_.size(); i++) {
        const int l = loadLocked(*devices_[i]);
        if (l < bestLoad) {
            bestLoad = l;
            best = i;
        }
    }
    return best;
}

// Waits (lock held on entry and exit) until every queue is empty and every device has been synchronised.
void ClDevicePool::globalDrainLocked(std::unique_lock<std::mutex>& lk) {
    for (auto& d : devices_) d->drainRequested = true;
    cvWork_.notify_all();
    cvIdle_.wait(lk, [&] {
        if (!running_) return true;
        if (!shared_.empty()) return false;
        for (auto& d : devices_)
            if (!d->priv.empty() || d->inFlight != 0 || d->drainRequested) return false;
        return true;
    });
}

void ClDevicePool::producerLoop() {
    std::unique_lock<std::mutex> lk(mu_);
    while (running_) {
        cvPools_.wait(lk, [&] { return !running_ || !pools_.empty(); });
        if (!running_) break;
        ClTaskPool pool = std::move(pools_.front());
        pools_.pop_front();
        producerBusy_ = true;
        int selected = -1;      // device of the current DEVICE_SELECT / SERIAL_MODE run, or -1
        bool serial = false;
        for (const ClTask& t : pool.tasks()) {
            if (!running_ || devices_.empty()) break;
            if (t.type & TASK_MESSAGE_GLOBAL_SYNCHRONIZATION_FIRST) globalDrainLocked(lk);
            if (t.type & (TASK_MESSAGE_DEVICE_SELECT_BEGIN | TASK_MESSAGE_SERIAL_MODE_BEGIN)) {
                selected = leastLoadedLocked();
                serial = (t.type & TASK_MESSAGE_SERIAL_MODE_BEGIN) != 0;
            }
            Item item{t, serial};
            if (t.type & TASK_MESSAGE_BROADCAST) {
                for (auto& d : devices_) d->priv.push_back(item);  // one copy per device (ClPipeline.cs:4264-4275)
            } else if (selected >= 0) {
                devices_[selected]->priv.push_back(item);
            } else if (type_ == DEVICE_ROUND_ROBIN) {
                devices_[rrCursor_++ % devices_.size()]->priv.push_back(item);
            } else {
                shared_.push_back(item);  // whoever is free first takes it
            }
            cvWork_.notify_all();
            if (t.type & (TASK_MESSAGE_DEVICE_SELECT_END | TASK_MESSAGE_SERIAL_MODE_END)) {
                selected = -1;
                serial = false;
            }
            if (t.type & TASK_MESSAGE_GLOBAL_SYNCHRONIZATION_LAST) globalDrainLocked(lk);
        }
        producerBusy_ = false;
        cvIdle_.notify_all();
    }
}

void ClDevicePool::consumerLoop(Device* d) {
    ClNumberCruncher& cr = *d->cruncher;
    std::unique_lock<std::mutex> lk(mu_);
    while (running_) {
        cvWork_.wait(lk, [&] { return !running_ || d->drainRequested || !d->priv.empty() || !shared_.empty(); });
        if (!running_) break;
        Item item;
        bool have = false;
        if (!d->priv.empty()) {
            item = std::move(d->priv.front());
            d->priv.pop_front();
            have = true;
        } else if (!shared_.empty()) {
            bool room = true;
            if (fineGrained_) {
                // throttle on markers still in flight so that a fast enqueuer cannot swallow the whole pool (:4899-4909)
                lk.unlock();
                room = cr.countMarkersRemaining() < deviceQueueLimit_;
                lk.lock();
            }
            if (room && !shared_.empty()) {
                item = std::move(shared_.front());
                shared_.pop_front();
                have = true;
            } else if (!room) {
                lk.unlock();
                std::this_thread::yield();
                lk.lock();
                continue;
            }
        }
        if (have) {
            d->inFlight++;
            lk.unlock();
            if (cr.errorCode() == 0) {
                if (item.task.type & TASK_MESSAGE_NO_COMPUTE) cr.setNoComputeMode(true);
                if (fineGrained_) {
                    if (!cr.enqueueMode()) cr.setEnqueueMode(true);
                    cr.setFineGrainedQueueControl(true);
                    if (!item.serial) cr.setEnqueueModeAsyncEnable(true);  // next queue, round-robin (:4946-4953)
                    item.task.compute(cr);
                    if (!item.serial) cr.setEnqueueModeAsyncEnable(false);
                    if (item.task.callback) d->callbacks.push_back(item.task.callback);
                } else {
                    item.task.compute(cr);  // blocking: the device is busy until the task is done → greedy balance
                    if (item.task.callback) item.task.callback();
                }
                if (item.task.type & TASK_MESSAGE_NO_COMPUTE) cr.setNoComputeMode(false);
            }
            lk.lock();
            d->consumed++;
            if (!fineGrained_) d->inFlight--;
            cvIdle_.notify_all();
            continue;
        }
        if (d->drainRequested) {
            // nothing left for this device: synchronise it and run the callbacks of everything that completed
            lk.unlock();
            if (cr.enqueueMode()) cr.setEnqueueMode(false);
            cr.sync();
            auto cbs = std::move(d->callbacks);
            d->callbacks.clear();
            for (auto& cb : cbs) cb();
            lk.lock();
            d->inFlight = 0;
            d->drainRequested = false;
            cvIdle_.notify_all();
        }
    }
    lk.unlock();
    if (cr.enqueueMode()) cr.setEnqueueMode(false);
    cr.sync();
}

double ClDevicePool::finish() {
    std::unique_lock<std::mutex> lk(mu_);
    cvIdle_.wait(lk, [&] { return !running_ || (pools_.empty() && !producerBusy_); });
    if (running_ && !devices_.empty()) globalDrainLocked(lk);
    double ms = 0.0;
    if (timing_) ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0_).count();
    timing_ = false;
    return ms;
}

void ClDevicePool::dispose() {
    {
        std::lock_guard<std::mutex> lk(mu_);
        if (!running_) return;
        running_ = false;
    }
    cvPools_.notify_all();
    cvWork_.notify_all();
    cvIdle_.notify_all();
    if (producer_.joinable()) producer_.join();
    for (auto& d : devices_)
        if (d->thread.joinable()) d->thread.join();
}

}  // namespace cekirdekler
