// cruncher.h — user-facing layer: ClArray (+flags), ClParameterGroup (nextParam chaining), ClNumberCruncher.
//
// C++ mirror of the reference's L4 API: ClNumberCruncher.cs:32-416, ClArray.cs:219-659 and
// :717-1907.  Same names, same argument meaning, same validation rules and the same "nothing throws, errors
// are counted and printed" behaviour (ClArray.cs:1605-1679), so tests written against this layer read like
// the reference's Tester.cs.  The C# generics/indexer sugar is left to the language binding above this layer.
#pragma once
#include "cores.h"

#include <memory>
#include <string>
#include <vector>

namespace cekirdekler {

// CSpaceArrays.ARR_* (CSpaceArrays.cs:48-78)
enum ArrType { ARR_FLOAT = 0, ARR_DOUBLE = 1, ARR_LONG = 2, ARR_INT = 3, ARR_UINT = 4, ARR_BYTE = 5, ARR_CHAR = 6 };
int arrTypeSize(int arrType);
int arrTypeToClType(int arrType);  // ARR_* → ClBuffer.SizeOf (CSpaceArrays.cs:290-317)

class ClNumberCruncher;

class ClArray {
  public:
    // FastArr-backed ("C++ array", ClArray.cs:889-958 fastArr=true): pinned, aligned native memory
    ClArray(int arrType, int64_t n, int alignmentBytes = 4096);
    // wraps caller-owned memory (the C# `T[]` case): never freed here; pageable unless the caller registered it
    ClArray(void* userMemory, int arrType, int64_t n);
    ~ClArray();
    ClArray(const ClArray&) = delete;
    ClArray& operator=(const ClArray&) = delete;

    void* data() const { return host_; }
    int64_t Length() const { return n_; }
    int arrType() const { return arrType_; }
    bool isFastArr() const { return fast_ != nullptr; }
    int elementSize() const { return arrTypeSize(arrType_); }

    // flag properties with the reference's mutual-exclusion rules (ClArray.cs:1742-1863)
    bool read() const { return read_; }
    void setRead(bool v) { if (!writeOnly_) read_ = v; }
    bool partialRead() const { return partialRead_; }
    void setPartialRead(bool v) { if (!writeOnly_) partialRead_ = v; }
    bool write() const { return write_; }
    void setWrite(bool v) { if (!readOnly_) write_ = v; }
    bool writeAll() const { return writeAll_; }
    void setWriteAll(bool v) { if (!readOnly_) writeAll_ = v; }
    bool readOnly() const { return readOnly_; }
    void setReadOnly(bool v);
    bool writeOnly() const { return writeOnly_; }
    void setWriteOnly(bool v);
    bool zeroCopy = false;
    int numberOfElementsPerWorkItem = 1;
    int alignmentBytes = 4096;

    uint32_t flagMask() const;        // the F_* bitmask equivalent of the flag string
    std::string flagString() const;   // " partial  read  write  all  ro  wo  zc " (ClArray.cs:1685-1703)
    ArrayRef ref() const;

  private:
    void* host_ = nullptr;
    ck_handle fast_ = nullptr;
    int64_t n_ = 0;
    int arrType_ = ARR_FLOAT;
    bool read_ = true, partialRead_ = false, write_ = true, writeAll_ = false, readOnly_ = false, writeOnly_ = false;
};

// a.nextParam(b, c): ordered kernel parameter list (ClArray.cs:219-500, :1365-1537)
class ClParameterGroup {
  public:
    ClParameterGroup() {}
    ClParameterGroup& nextParam(ClArray* a) { arrays.push_back(a); return *this; }
    std::vector<ClArray*> arrays;
    // ClArray.cs:543-659 / :1605-1736.  Returns 0 when the compute was dispatched, else the validation error number.
    int compute(ClNumberCruncher& cruncher, int computeId, const std::string& kernelNamesString, int globalRange, int localRange = 256,
                int ofsetGlobalRange = 0, bool pipeline = false, bool pipelineType = Cores::PIPELINE_EVENT, int pipelineBlobs = 4,
                const std::vector<std::string>* readWritesReady = nullptr, const std::vector<int>* elementsPerItemReady = nullptr);
};

class ClNumberCruncher {
  public:
    // ClNumberCruncher.cs:313-349 (explicit device list).  `streamingEnabled` is the ClDevices GPU_STREAM switch.
    ClNumberCruncher(const std::vector<int>& deviceOrdinals, const std::string& kernelString, bool noPipelining = false,
                     int computeQueueConcurrency = 16, bool streamingEnabled = false);
    ~ClNumberCruncher();

    static std::vector<std::string> extractKernelNames(const std::string& kernelString);  // ClNumberCruncher.cs:219-228

    int errorCode() const { return errorNotification_; }
    const std::string& errorMessage() const { return errorMessage_; }
    int numberOfErrorsHappened = 0;
    Cores* numberCruncher() { return cores_.get(); }
    int numberOfDevices() const { return cores_ ? cores_->numberOfDevices() : 0; }

    // properties (ClNumberCruncher.cs:66-187)
    int repeatCount = 1;
    std::string repeatKernelName;
    bool performanceFeed = false;
    void setEnqueueMode(bool v) { if (cores_) cores_->setEnqueueMode(v); }
    bool enqueueMode() const { return cores_ && cores_->enqueueMode(); }
    void setEnqueueModeAsyncEnable(bool v) { if (cores_) cores_->setEnqueueModeAsyncEnable(v); }
    void setNoComputeMode(bool v) { if (cores_) cores_->noComputeMode = v; }
    void setFineGrainedQueueControl(bool v) { if (cores_) cores_->fineGrainedQueueControl = v; }
    void setSmoothLoadBalancer(bool v) { if (cores_) cores_->smoothLoadBalancer = v; }
    int countMarkersRemaining() { return cores_ ? cores_->countMarkers() - cores_->countMarkerCallbacks() : 0; }
    int countMarkersReached() { return cores_ ? cores_->countMarkerCallbacks() : 0; }
    void flushLastUsedCommandQueue() { if (cores_) cores_->flushLastUsedCommandQueue(); }
    void sync();  // empty in the reference (quirk Q6); here: drain every queue of every device
    std::string lastComputePerformanceReport;

  private:
    std::unique_ptr<Cores> cores_;
    int errorNotification_ = 0;
    std::string errorMessage_;
};

}  // namespace cekirdekler
