// pool.h — task pool / device pool: greedy scheduling of frozen compute() calls over a set of devices.
//
// B200-native restatement of the reference's Pipeline.Pool namespace (ClPipeline.cs:3241-5078):
//   ClTask        a frozen ClArray/ClParameterGroup.compute() call (:3331-3520)
//   ClTaskPool    an ordered list of tasks (:3607-3783)
//   ClDevicePool  a producer thread + one consumer thread per device (:3872-5047).  The producer walks the pool in order
//                 and honours the control bits of ClTaskType (:3247-3321): GLOBAL_SYNCHRONIZATION_FIRST/LAST (barriers over
//                 all devices), DEVICE_SELECT_BEGIN/END (a run of tasks pinned to the least-loaded device, :4100-4127),
//                 SERIAL_MODE_BEGIN/END (same device AND same queue), BROADCAST (one copy per device, :4264-4275).
//                 Every consumer pulls work greedily from the shared FIFO (DEVICE_COMPUTE_AT_WILL) and issues it in
//                 enqueue mode, round-robin over up to `maxQueuesPerDevice` streams (:4946-4953), throttled by the number
//                 of markers still in flight when fine-grained control is on (:4899-4909).
// Differences: synchronisation uses one mutex + condition variables instead of Monitor ping-pong over message pipes, and
// DEVICE_ROUND_ROBIN (declared but not implemented in the reference, :3806) deals tasks to devices in turn.
#pragma once
#include "cruncher.h"

#include <atomic>
#include <condition_variable>
#include <deque>
#include <functional>
#include <memory>
#include <thread>

namespace cekirdekler {

enum ClTaskType : int {  // ClPipeline.cs:3247-3321
    TASK_MESSAGE_DEFAULT = 0,
    TASK_MESSAGE_DEVICE_SELECT_BEGIN = 1,
    TASK_MESSAGE_DEVICE_SELECT_END = 2,
    TASK_MESSAGE_GLOBAL_SYNCHRONIZATION_FIRST = 4,
    TASK_MESSAGE_GLOBAL_SYNCHRONIZATION_LAST = 8,
    TASK_MESSAGE_BROADCAST = 16,
    TASK_MESSAGE_NO_COMPUTE = 32,
    TASK_MESSAGE_SERIAL_MODE_BEGIN = 64,
    TASK_MESSAGE_SERIAL_MODE_END = 128,
    TASK_MESSAGE_LAST = 256,
};

struct ClTask {  // ClPipeline.cs:3331-3520; built by ClArray.task()/ClParameterGroup.task() (ClArray.cs:515-521, :1552-1583)
    int type = TASK_MESSAGE_DEFAULT;
    std::vector<ClArray*> arrays;          // the parameter list
    std::vector<std::string> readWrite;    // flag strings frozen at task() time
    std::vector<int> elementsPerItem;
    int computeId = 0;
    std::string kernelNamesString;
    int globalRange = 0, localRange = 256, ofsetGlobalRange = 0;
    bool pipeline = false, pipelineType = Cores::PIPELINE_EVENT;
    int pipelineBlobs = 4;
    std::function<void()> callback;        // setCallBack (:3481): runs once, after the task has completed on its device
    int compute(ClNumberCruncher& cruncher) const;
};

class ClTaskPool {  // ClPipeline.cs:3607-3783
  public:
    void feed(const ClTask& task) { tasks_.push_back(task); }
    void reset() { counter_ = 0; }
    const ClTask* nextTask() { return counter_ < tasks_.size() ? &tasks_[counter_++] : nullptr; }
    int remainingTaskGroupsOrTasks() const { return (int)(tasks_.size() - counter_); }
    int numTasks() const { return (int)tasks_.size(); }
    const std::vector<ClTask>& tasks() const { return tasks_; }

  private:
    std::vector<ClTask> tasks_;
    size_t counter_ = 0;
};

enum ClDevicePoolType : int { DEVICE_COMPUTE_AT_WILL = 0, DEVICE_ROUND_ROBIN = 1 };  // ClPipeline.cs:3802-3808

class ClDevicePool {  // ClPipeline.cs:3872-4700
  public:
    ClDevicePool(ClDevicePoolType poolType, const std::string& kernelCodeToCompile, bool fineGrainedQueueControl = false,
                 int maxQueuesPerDevice = 3);
    ~ClDevicePool();
    void addDevice(int deviceOrdinal);                 // :4340  (one consumer thread + one ClNumberCruncher per device)
    void enqueueTaskPool(const ClTaskPool& pool);      // :4400  (the pool is copied; the caller may reuse it)
    double finish();                                   // :4470  waits for everything enqueued so far; elapsed ms
    void dispose();                                    // :4527
    int numberOfDevices() const { return (int)devices_.size(); }
    int errorCode() const;
    std::string errorMessage() const;
    std::vector<int> tasksPerDevice() const;           // how many tasks each device consumed (the greedy split)

  private:
    struct Item {
        ClTask task;
        bool serial = false;
    };
    struct Device {
        int ordinal = 0;
        std::unique_ptr<ClNumberCruncher> cruncher;
        std::deque<Item> priv;                         // tasks addressed to this device (select / serial / broadcast)
        std::thread thread;
        int consumed = 0;
        int inFlight = 0;                              // tasks issued since the last drain
        bool drainRequested = false;
        std::vector<std::function<void()>> callbacks;  // run at the next drain
    };
    void producerLoop();
    void consumerLoop(Device* d);
    void globalDrainLocked(std::unique_lock<std::mutex>& lk);  // all queues empty, all devices synchronised
    int leastLoadedLocked() const;
    int loadLocked(const Device& d) const;

    ClDevicePoolType type_;
    std::string kernelCode_;
    bool fineGrained_;
    int maxQueues_;
    int deviceQueueLimit_;
    std::vector<std::unique_ptr<Device>> devices_;
    mutable std::mutex mu_;
    std::condition_variable cvWork_, cvIdle_, cvPools_;
    std::deque<ClTaskPool> pools_;
    std::deque<Item> shared_;
    bool producerBusy_ = false;
    bool running_ = true;
    int rrCursor_ = 0;
    std::thread producer_;
    std::chrono::steady_clock::time_point t0_;
    bool timing_ = false;
};

}  // namespace cekirdekler
