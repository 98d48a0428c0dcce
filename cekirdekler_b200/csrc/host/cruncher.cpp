// cruncher.cpp — see cruncher.h.  Reference: ClNumberCruncher.cs:199-349, ClArray.cs:543-659, :1605-1736,
// :1742-1863, CSpaceArrays.cs:290-320.
#include "cruncher.h"

#include <cstdio>
#include <regex>

namespace cekirdekler {

int arrTypeSize(int t) {
    static int table[20] = {-1};
    if (table[0] == -1) sizeOf(table);  // "loading type sizes from C space" (CSpaceArrays.cs:88-91)
    return (t >= 0 && t < 20) ? table[t] : 0;
}

int arrTypeToClType(int t) {
    switch (t) {
        case ARR_FLOAT: return 0;   // cl_float
        case ARR_DOUBLE: return 1;  // cl_double
        case ARR_INT: return 2;     // cl_int
        case ARR_LONG: return 3;    // cl_long
        case ARR_CHAR: return 4;    // cl_half (2-byte C# char)
        case ARR_BYTE: return 5;    // cl_char
        case ARR_UINT: return 6;    // cl_uint
        default: return 0;
    }
}

// ------------------------------------------------------------------ ClArray
ClArray::ClArray(int arrType, int64_t n, int alignment) : n_(n), arrType_(arrType) {
    alignmentBytes = alignment;
    fast_ = createArray((int32_t)n, alignment, arrType);
    host_ = fast_ ? alignedArrHead(fast_) : nullptr;
}
ClArray::ClArray(void* userMemory, int arrType, int64_t n) : host_(userMemory), n_(n), arrType_(arrType) {}
ClArray::~ClArray() {
    if (fast_) deleteArray(fast_);
}
void ClArray::setReadOnly(bool v) {
    if (v && !writeOnly_) {
        write_ = false;
        writeAll_ = false;
        writeOnly_ = false;
        readOnly_ = true;
    } else if (!v) {
        readOnly_ = false;
    }
}
void ClArray::setWriteOnly(bool v) {
    if (v && !readOnly_) {
        read_ = false;
        partialRead_ = false;
        readOnly_ = false;
        writeOnly_ = true;
    } else if (!v) {
        writeOnly_ = false;
    }
}
uint32_t ClArray::flagMask() const {
    uint32_t f = 0;
    if (partialRead_) f |= F_PARTIAL;
    if (read_) f |= F_READ;
    if (write_) f |= F_WRITE;
    if (writeAll_) f |= F_ALL;
    if (readOnly_) f |= F_RO;
    if (writeOnly_) f |= F_WO;
    if (zeroCopy) f |= F_ZC;
    return f;
}
std::string ClArray::flagString() const {
    std::string s;
    if (partialRead_) s += " partial ";
    if (read_) s += " read ";
    if (write_) s += " write ";
    if (writeAll_) s += " all ";
    if (readOnly_) s += " ro ";
    if (writeOnly_) s += " wo ";
    if (zeroCopy) s += " zc ";
    return s;
}
ArrayRef ClArray::ref() const {
    ArrayRef r;
    r.host = host_;
    r.length = n_;
    r.clType = arrTypeToClType(arrType_);
    r.fastArr = isFastArr();
    r.flags = flagMask();
    r.elemsPerWorkItem = numberOfElementsPerWorkItem;
    // cache key: FastArr-backed arrays are keyed by object, wrapped caller memory by its address (the C# array object)
    r.identity = fast_ ? (const void*)this : (const void*)host_;
    return r;
}

// ------------------------------------------------------------------ ClNumberCruncher
std::vector<std::string> ClNumberCruncher::extractKernelNames(const std::string& src) {
    // "kernel _ void _ name (" — also matches inside "__kernel" exactly as the reference's regex does
    std::vector<std::string> names;
    static const std::regex re("kernel[\\s]+void[\\s]+([a-zA-Z\\d_]+)");
    for (auto it = std::sregex_iterator(src.begin(), src.end(), re); it != std::sregex_iterator(); ++it) {
        std::string n = (*it)[1].str();
        bool dup = false;
        for (auto& e : names) dup = dup || (e == n);
        if (!dup) names.push_back(n);
    }
    return names;
}

ClNumberCruncher::ClNumberCruncher(const std::vector<int>& deviceOrdinals, const std::string& kernelString, bool noPipelining,
                                   int computeQueueConcurrency, bool streamingEnabled) {
    auto names = extractKernelNames(kernelString);
    if (names.empty()) {
        errorMessage_ = "Error: no kernel definitions are found in string. Kernel string: \n" + kernelString;
        fprintf(stderr, "%s\n", errorMessage_.c_str());
        errorNotification_ = 1;
        return;
    }
    cores_.reset(new Cores(deviceOrdinals, kernelString, names, computeQueueConcurrency, noPipelining, streamingEnabled));
    if (cores_->errorCode() != 0) {
        errorMessage_ = cores_->errorMessage();
        fprintf(stderr, "%s\n", errorMessage_.c_str());
        errorNotification_ = cores_->errorCode();
        cores_.reset();  // the reference disposes the cores; the cruncher stays inert (ClNumberCruncher.cs:340-347)
        numberOfErrorsHappened++;
    }
}
ClNumberCruncher::~ClNumberCruncher() {}

void ClNumberCruncher::sync() {
    if (!cores_) return;
    for (int i = 0; i < cores_->numberOfDevices(); i++) cores_->worker(i).finishAll();
}

// ------------------------------------------------------------------ ClParameterGroup.compute
static std::vector<std::string> splitKernelNames(const std::string& s) {
    // separators of ClArray.cs:1681: space , ; - newline @
    std::vector<std::string> out;
    std::string cur;
    for (char c : s) {
        if (c == ' ' || c == ',' || c == ';' || c == '-' || c == '\n' || c == '@' || c == '\r' || c == '\t') {
            if (!cur.empty()) out.push_back(cur);
            cur.clear();
        } else {
            cur.push_back(c);
        }
    }
    if (!cur.empty()) out.push_back(cur);
    return out;
}

int ClParameterGroup::compute(ClNumberCruncher& cr, int computeId, const std::string& kernelNamesString, int globalRange, int localRange,
                              int ofsetGlobalRange, bool pipeline, bool pipelineType, int pipelineBlobs,
                              const std::vector<std::string>* readWritesReady, const std::vector<int>* elementsPerItemReady) {
    auto complain = [&](int code, const std::string& msg) {
        fprintf(stderr, "\n%s\n\n", msg.c_str());
        cr.numberOfErrorsHappened++;
        return code;
    };
    if (cr.numberOfErrorsHappened > 0) {
        fprintf(stderr, "\nError: there has been %d error(s) before, cannot compute.\n\n", cr.numberOfErrorsHappened);
        return 1;
    }
    if (cr.errorCode() != 0 || !cr.numberCruncher()) {
        fprintf(stderr, "Number-cruncher C99-device compiling error:\n%s\n", cr.errorMessage().c_str());
        cr.numberOfErrorsHappened++;
        return 2;
    }
    Cores& cores = *cr.numberCruncher();
    const long long D = cores.spmdWorld() > 0 ? cores.spmdWorld() : cores.numberOfDevices();
    if (localRange <= 0 || (globalRange % localRange) != 0)
        return complain(3, "Work-size error: global range(" + std::to_string(globalRange) +
                               ") is not an integer multiple of local range. Global range has to be exact multiple of local range(" +
                               std::to_string(localRange) + ").  This is obliged by OpenCL rules.");
    if (pipeline) {
        if (pipelineBlobs <= 0 || (globalRange % (localRange * pipelineBlobs)) != 0)
            return complain(4, "Work-size error: global range(" + std::to_string(globalRange) +
                                   ") is not an integer multiple of (local range)*(number of pipeline blobs)=(" +
                                   std::to_string((long long)localRange * pipelineBlobs) + ").");
        // the event-driven pipeline runs two lanes of blobs/2 chunks each (Cores.cs:1256-1366): with an odd count the last
        // chunk of each lane would be neither uploaded, computed nor downloaded (the reference only documents "multiple of 4")
        if (pipelineType == Cores::PIPELINE_EVENT && (pipelineBlobs % 2) != 0)
            return complain(11, "Work-size error: the event-driven pipeline needs an even number of pipeline blobs (" +
                                    std::to_string(pipelineBlobs) + " given); use a multiple of 4, or the driver-controlled pipeline.");
        if (globalRange < (long long)localRange * D * pipelineBlobs)
            return complain(6, "Work-size error: global range(" + std::to_string(globalRange) +
                                   ") must be equal to or greater than (number of selected devices)*(local worksize)*(number of pipeline blobs)=(" +
                                   std::to_string((long long)localRange * D * pipelineBlobs) + ") if pipelining is enabled.");
    } else if (globalRange < (long long)localRange * D) {
        return complain(5, "Work-size error: global work size(" + std::to_string(globalRange) +
                               ") must be equal to or greater than (number of selected devices)*(local worksize)=(" +
                               std::to_string((long long)localRange * D) + ")");
    }
    std::vector<ArrayRef> refs;
    for (size_t i = 0; i < arrays.size(); i++) {
        ArrayRef r = arrays[i]->ref();
        if (readWritesReady && i < readWritesReady->size()) r.flags = parseFlags((*readWritesReady)[i].c_str());
        if (elementsPerItemReady && i < elementsPerItemReady->size()) r.elemsPerWorkItem = (*elementsPerItemReady)[i];
        // only sliced transfers constrain the length; whole-array copies move Length elements whatever the range
        if ((r.flags & F_PARTIAL) || ((r.flags & F_WRITE) && !(r.flags & F_ALL))) {
            if (r.length < (long long)globalRange * r.elemsPerWorkItem)
                return complain(7, "Array-size error: (global range)*(number of array elements per work item)=(" +
                                       std::to_string((long long)globalRange * r.elemsPerWorkItem) +
                                       ") must be equal to or less than array length (" + std::to_string(r.length) + ").");
        }
        if (!r.host) return complain(8, "Array error: parameter " + std::to_string(i) + " has no host memory.");
        refs.push_back(r);
    }
    auto kernels = splitKernelNames(kernelNamesString);
    if (kernels.empty()) return complain(9, "Kernel-name error: no kernel names given.");
    int launchErrorsBefore = 0;
    for (int i = 0; i < cores.numberOfDevices(); i++) launchErrorsBefore += cores.worker(i).launchErrors();
    cores.compute(kernels, cr.repeatCount, cr.repeatCount > 1 ? cr.repeatKernelName : std::string(), refs, globalRange, computeId,
                  ofsetGlobalRange, pipeline, pipelineBlobs, pipelineType, localRange);
    int launchErrorsAfter = 0;
    for (int i = 0; i < cores.numberOfDevices(); i++) launchErrorsAfter += cores.worker(i).launchErrors();
    if (launchErrorsAfter != launchErrorsBefore) return 10;  // a kernel could not be launched (message already printed)
    if (cr.performanceFeed) {
        cr.lastComputePerformanceReport = cores.performanceReport(computeId);
        fprintf(stdout, "%s\n", cr.lastComputePerformanceReport.c_str());
    }
    return 0;
}

}  // namespace cekirdekler
