// pipeline.h — device-to-device stage pipeline (ClPipeline / ClPipelineStage / ClPipelineStageBuffer).
//
// B200-native restatement of the reference's Pipeline namespace (ClPipeline.cs:41-2356).  Each stage owns a
// ClNumberCruncher on its own device(s) and double-buffered inputs/outputs (`buf` / `bufDuplicate`); one
// pushData() runs every stage on its `buf`s while, concurrently, every stage's previous results (`bufDuplicate`)
// move to the next stage's input duplicate; then all pairs swap (ClPipeline.cs:49-125, :2177-2208).
// Results therefore appear after 2S-2 .. 2S pushes exactly as in the reference (:114-124).
//
// What changes is where the bytes travel.  The reference downloads every stage's outputs to host arrays, memcpy's them
// into the next stage's host arrays and uploads them again (ClPipeline.cs:1422-1574): two host-link crossings and a
// host copy per hand-off.  Here (Forwarding::PeerToPeer, the default) the stage buffers stay in HBM: kernels run on
// device-resident buffers with no transfers, and a hand-off is one ckCopyBufferPeer over NVLink from stage k's output
// duplicate into stage k+1's input duplicate.  Only the two ends touch the host: pushData's `data` arrays are uploaded
// into stage 0's input duplicates, and the last stage's output duplicates are downloaded into `popResultsHere`.
// Forwarding::Host keeps the reference's data path (for comparison and for parity tests).
#pragma once
#include "cruncher.h"

#include <memory>
#include <string>
#include <vector>

namespace cekirdekler {

enum class Forwarding { PeerToPeer = 0, Host = 1 };

class ClPipelineStageBuffer {  // ClPipeline.cs:1886-2356
  public:
    ClPipelineStageBuffer(int arrType, int64_t n, int elementsPerWorkItem);
    ClArray* buf() const { return cur_; }
    ClArray* bufDuplicate() const { return dup_; }
    void switchBuffers() { std::swap(cur_, dup_); }  // ClPipeline.cs:2177-2208
    int64_t length() const { return a_->Length(); }
    int arrType() const { return a_->arrType(); }

  private:
    std::unique_ptr<ClArray> a_, b_;
    ClArray* cur_;
    ClArray* dup_;
};

class ClPipeline;

class ClPipelineStage {  // ClPipeline.cs:146-1873
  public:
    explicit ClPipelineStage(bool debugLog = false) : debug_(debugLog) {}
    ~ClPipelineStage();

    void addDevices(const std::vector<int>& deviceOrdinals) { devices_ = deviceOrdinals; }                       // :1731
    void addKernels(const std::string& kernels, const std::string& kernelNames, const std::vector<int>& globalRanges,
                    const std::vector<int>& localRanges);                                                          // :1744
    void initializerKernel(const std::string& kernelNames, const std::vector<int>& globalRanges,
                           const std::vector<int>& localRanges);                                                   // :1678
    ClPipelineStageBuffer* addInputBuffer(int arrType, int64_t n, int elementsPerWorkItem = 1);                   // :1810-1843
    ClPipelineStageBuffer* addOutputBuffer(int arrType, int64_t n, int elementsPerWorkItem = 1);                  // :1845-1878
    ClPipelineStageBuffer* addHiddenBuffer(int arrType, int64_t n, int elementsPerWorkItem = 1);                  // :1777-1808
    void prependToStage(ClPipelineStage* stage);  // this stage runs BEFORE `stage`                                // :1704
    void appendToStage(ClPipelineStage* stage);   // this stage runs AFTER `stage`                                 // :1717
    ClPipeline* makePipeline(Forwarding mode = Forwarding::PeerToPeer);                                            // :1582
    bool enqueueMode = false;

    int errorCode() const { return cruncher_ ? cruncher_->errorCode() : initError_; }
    std::string errorMessage() const { return cruncher_ ? cruncher_->errorMessage() : initMessage_; }
    double elapsedTime = 0.0;  // ms of the last run() (device span)

  private:
    friend class ClPipeline;
    struct KernelParameters { std::string name; int globalRange; int localRange; };
    bool init(Forwarding mode);
    void run(bool initializerKernels, Forwarding mode);
    std::vector<ArrayRef> parameterRefs(Forwarding mode) const;  // inputs, hidden, outputs (ClPipeline.cs:281-370)

    std::vector<int> devices_;
    std::string kernelsToCompile_;
    std::vector<KernelParameters> kernels_, initKernels_;
    std::vector<std::unique_ptr<ClPipelineStageBuffer>> inputs_, outputs_, hiddens_;
    ClPipelineStage* previous_ = nullptr;
    ClPipelineStage* next_ = nullptr;
    std::unique_ptr<ClNumberCruncher> cruncher_;
    bool initComplete_ = false;
    bool initKernelsDone_ = false;
    int initError_ = 0;
    std::string initMessage_;
    bool debug_;
    ck_handle forwardQueue_ = nullptr;  // a queue on this stage's device used for incoming hand-offs
    // events of the chained (peer-to-peer) pushData, two deep (push p uses [p % 2], waits on [(p - 1) % 2]):
    // this stage's kernels / the transfer that filled its input duplicates / the download that read its output duplicates
    ck_handle compDone_[2] = {nullptr, nullptr}, inDone_[2] = {nullptr, nullptr}, outDone_[2] = {nullptr, nullptr};
    ck_handle forwardCtx_ = nullptr, forwardDev_ = nullptr, forwardPlat_ = nullptr;
};

class ClPipeline {  // ClPipeline.cs:41-139
  public:
    ~ClPipeline();
    // data / popResultsHere: host arrays (one per input of the first stage / output of the last stage), or empty.
    // Returns true once the results delivered by this push are valid (ClPipeline.cs:114-124).
    bool pushData(const std::vector<void*>& data = {}, const std::vector<void*>& popResultsHere = {});
    int numberOfStages() const { return (int)stages_.size(); }
    int counter = 0;
    int errorCode() const;
    std::string errorMessage() const;
    Forwarding forwarding() const { return mode_; }

  private:
    friend class ClPipelineStage;
    bool pushChained(const std::vector<void*>& data, const std::vector<void*>& popResultsHere);
    bool finishPush(bool hasData, bool hasPop);
    std::vector<ClPipelineStage*> stages_;
    Forwarding mode_ = Forwarding::PeerToPeer;
    bool chained_ = false;   // every stage on one device, peer-to-peer forwarding: pushes are ordered by events, not by a host join
};

}  // namespace cekirdekler

// ====================================================================================================================
// SingleGPUPipeline: N kernel stages overlapped on ONE device (ClPipeline.cs:2357-3240)
// ====================================================================================================================
namespace cekirdekler {

enum class DevicePipelineArrayType { INPUT = 0, OUTPUT = 1, INTERNAL = 2, TRANSITION = 3 };  // ClPipeline.cs:3178-3184

// An array bound to one or two stages.  INPUT/OUTPUT wrap the caller's host array (uploaded before / downloaded after the
// stage's kernels, on the stage's own queue); INTERNAL lives only in HBM; TRANSITION is shared by two consecutive stages
// and double-buffered in HBM: the producer writes one copy while the consumer reads the copy of the previous feed().
class DevicePipelineArray {  // ClPipeline.cs:3186-3240
  public:
    DevicePipelineArray(DevicePipelineArrayType type, ClArray* userArray);                         // INPUT / OUTPUT
    DevicePipelineArray(DevicePipelineArrayType type, int arrType, int64_t n, int elementsPerWorkItem = 1);  // INTERNAL / TRANSITION
    DevicePipelineArrayType type;
    ClArray* version(int v) const { return (v & 1) && b_ ? b_.get() : (user_ ? user_ : a_.get()); }
    int producerStage = -1;  // first stage that bound a TRANSITION array

  private:
    ClArray* user_ = nullptr;
    std::unique_ptr<ClArray> a_, b_;
};

class DevicePipelineStage {  // ClPipeline.cs:2636-3176
  public:
    DevicePipelineStage(const std::string& kernelNames, int globalRange, int localRange)
        : kernelNames(kernelNames), globalRange(globalRange), localRange(localRange) {}
    void bindArray(DevicePipelineArray* a) { arrays.push_back(a); }                                // :3106
    bool stopHostDeviceTransmission = false;                                                       // :2638
    std::string kernelNames;
    int globalRange, localRange;
    std::vector<DevicePipelineArray*> arrays;
};

class DevicePipeline {  // ClPipeline.cs:2359-2634
  public:
    DevicePipeline(int deviceOrdinal, const std::string& kernelCodes, int computeQueueConcurrency = 16);
    void addStage(DevicePipelineStage* stage);                                                     // :2381
    void feed();                                                                                   // :2570
    void feedAsyncBegin();                                                                         // :2603
    void feedAsyncEnd();                                                                           // :2616
    void enableSerialMode() { serialMode_ = true; }                                                // :2621
    void enableParallelMode() { serialMode_ = false; }                                             // :2625
    int errorCode() const { return cruncher_ ? cruncher_->errorCode() : 0; }
    std::string errorMessage() const { return cruncher_ ? cruncher_->errorMessage() : std::string(); }
    int runCounter() const { return runCounter_; }
    // feeds between handing an input in and its result appearing in the OUTPUT array: 0 in serial mode, stages-1 in parallel mode
    int latency() const { return serialMode_ ? 0 : (int)stages_.size() - 1; }

  private:
    void enqueueAll();
    int device_;
    std::string code_;
    int concurrency_;
    std::vector<DevicePipelineStage*> stages_;
    std::unique_ptr<ClNumberCruncher> cruncher_;
    bool serialMode_ = false;
    int runCounter_ = 0;
    bool open_ = false;
};

}  // namespace cekirdekler
