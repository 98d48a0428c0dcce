// pipeline.cpp — see pipeline.h.  Reference: ClPipeline.cs:41-139 (pushData), :146-603 (stage run), :624-1575
// (forwardResults), :1582-1873 (stage building), :1886-2356 (stage buffers).
#include "pipeline.h"

#include <cstdio>
#include <cstring>

namespace cekirdekler {

// ------------------------------------------------------------------ stage buffers
ClPipelineStageBuffer::ClPipelineStageBuffer(int arrType, int64_t n, int elementsPerWorkItem)
    : a_(new ClArray(arrType, n)), b_(new ClArray(arrType, n)), cur_(a_.get()), dup_(b_.get()) {
    a_->numberOfElementsPerWorkItem = elementsPerWorkItem;
    b_->numberOfElementsPerWorkItem = elementsPerWorkItem;
}

// ------------------------------------------------------------------ stage building
static std::vector<std::string> splitNames(const std::string& s) {
    std::vector<std::string> out;
    std::string cur;
    for (char c : s) {
        if (c == ' ' || c == ',' || c == ';' || c == '-' || c == '\n' || c == '@' || c == '\t' || c == '\r') {
            if (!cur.empty()) out.push_back(cur);
            cur.clear();
        } else {
            cur.push_back(c);
        }
    }
    if (!cur.empty()) out.push_back(cur);
    return out;
}

void ClPipelineStage::addKernels(const std::string& kernels, const std::string& kernelNames, const std::vector<int>& globalRanges,
                                 const std::vector<int>& localRanges) {
    kernelsToCompile_ += "\n" + kernels;
    auto names = splitNames(kernelNames);
    for (size_t i = 0; i < names.size(); i++) {
        const int g = i < globalRanges.size() ? globalRanges[i] : (globalRanges.empty() ? 0 : globalRanges.back());
        const int l = i < localRanges.size() ? localRanges[i] : (localRanges.empty() ? 256 : localRanges.back());
        kernels_.push_back({names[i], g, l});
    }
}

void ClPipelineStage::initializerKernel(const std::string& kernelNames, const std::vector<int>& globalRanges,
                                        const std::vector<int>& localRanges) {
    auto names = splitNames(kernelNames);
    for (size_t i = 0; i < names.size(); i++) {
        const int g = i < globalRanges.size() ? globalRanges[i] : (globalRanges.empty() ? 0 : globalRanges.back());
        const int l = i < localRanges.size() ? localRanges[i] : (localRanges.empty() ? 256 : localRanges.back());
        initKernels_.push_back({names[i], g, l});
    }
}

ClPipelineStageBuffer* ClPipelineStage::addInputBuffer(int arrType, int64_t n, int e) {
    inputs_.emplace_back(new ClPipelineStageBuffer(arrType, n, e));
    return inputs_.back().get();
}
ClPipelineStageBuffer* ClPipelineStage::addOutputBuffer(int arrType, int64_t n, int e) {
    outputs_.emplace_back(new ClPipelineStageBuffer(arrType, n, e));
    return outputs_.back().get();
}
ClPipelineStageBuffer* ClPipelineStage::addHiddenBuffer(int arrType, int64_t n, int e) {
    hiddens_.emplace_back(new ClPipelineStageBuffer(arrType, n, e));
    return hiddens_.back().get();
}

void ClPipelineStage::prependToStage(ClPipelineStage* stage) {
    next_ = stage;
    if (stage) stage->previous_ = this;
}
void ClPipelineStage::appendToStage(ClPipelineStage* stage) {
    previous_ = stage;
    if (stage) stage->next_ = this;
}

ClPipelineStage::~ClPipelineStage() {
    for (int b = 0; b < 2; b++)
        for (ck_handle e : {compDone_[b], inDone_[b], outDone_[b]})
            if (e) deleteEvent(e);
    if (forwardQueue_) {
        finish(forwardQueue_);
        deleteCommandQueue(forwardQueue_);
    }
    if (forwardCtx_) deleteContext(forwardCtx_);
    if (forwardDev_) deleteDevice(forwardDev_);
    if (forwardPlat_) deletePlatform(forwardPlat_);
}

ClPipeline* ClPipelineStage::makePipeline(Forwarding mode) {
    // walk back to the first stage, then forward: one stage per layer, linear chains only (ClPipeline.cs:1596-1600)
    ClPipelineStage* first = this;
    while (first->previous_) first = first->previous_;
    auto* p = new ClPipeline();
    for (ClPipelineStage* s = first; s; s = s->next_) {
        p->stages_.push_back(s);
        if (s->devices_.size() != 1) mode = Forwarding::Host;  // a multi-device stage splits its buffers: keep the host path
    }
    p->mode_ = mode;
    for (auto* s : p->stages_) s->init(mode);
    const char* env = getenv("CK_PIPELINE_CHAINED");
    p->chained_ = mode == Forwarding::PeerToPeer && !(env && env[0] == '0');
    if (p->chained_)
        for (auto* s : p->stages_)
            for (int b = 0; b < 2; b++) {
                if (!s->compDone_[b]) s->compDone_[b] = createEvent();
                if (!s->inDone_[b]) s->inDone_[b] = createEvent();
                if (!s->outDone_[b]) s->outDone_[b] = createEvent();
            }
    return p;
}

bool ClPipelineStage::init(Forwarding mode) {
    if (initComplete_) return errorCode() == 0;
    initComplete_ = true;
    if (devices_.empty()) {
        initError_ = 1;
        initMessage_ = "pipeline stage has no devices";
        return false;
    }
    // "can't enable driver-pipelining but can have more device-pipeline-stages": noPipelining = true (ClPipeline.cs:231)
    cruncher_.reset(new ClNumberCruncher(devices_, kernelsToCompile_, true));
    if (cruncher_->errorCode() != 0) return false;
    // a queue of our own on the stage's (first) device for incoming hand-offs and end-of-pipeline transfers
    ck_handle list = platformList();
    forwardPlat_ = createPlatform(list, 0);
    deletePlatformList(list);
    forwardDev_ = createDevice(forwardPlat_, CODE_GPU(), devices_[0]);
    forwardCtx_ = forwardDev_ ? createContext(forwardPlat_, forwardDev_) : nullptr;
    forwardQueue_ = forwardCtx_ ? createCommandQueue(forwardCtx_, forwardDev_, 0) : nullptr;
    if (!forwardQueue_) {
        initError_ = 1;
        initMessage_ = std::string("pipeline stage: cannot create the hand-off queue: ") + ckGetLastError();
        return false;
    }
    if (mode == Forwarding::PeerToPeer) {
        // The reference's stage arrays start zeroed on the host and are uploaded; device-resident stage buffers must start
        // zeroed too (cudaMalloc does not clear), or the warm-up pushes would compute on whatever the memory held before.
        Worker& w = cruncher_->numberCruncher()->worker(0);
        for (auto* group : {&inputs_, &hiddens_, &outputs_})
            for (auto& sb : *group)
                for (ClArray* a : {sb->buf(), sb->bufDuplicate()}) {
                    ck_handle b = w.buffer(a->ref());
                    if (b) ckZeroBuffer(forwardQueue_, b);
                }
        finish(forwardQueue_);
    }
    if (!initKernels_.empty()) run(true, mode);
    return true;
}

std::vector<ArrayRef> ClPipelineStage::parameterRefs(Forwarding mode) const {
    // parameter order: inputs, hidden, outputs (ClPipeline.cs:281-370); flags as :236-270
    std::vector<ArrayRef> refs;
    for (auto& b : inputs_) {
        ArrayRef r = b->buf()->ref();
        r.flags = (mode == Forwarding::Host) ? F_READ : 0u;  // P2P: the data is already in HBM (arrived by hand-off)
        refs.push_back(r);
    }
    for (auto& b : hiddens_) {
        ArrayRef r = b->buf()->ref();
        r.flags = 0u;
        refs.push_back(r);
    }
    for (auto& b : outputs_) {
        ArrayRef r = b->buf()->ref();
        r.flags = (mode == Forwarding::Host) ? F_WRITE : 0u;
        refs.push_back(r);
    }
    return refs;
}

void ClPipelineStage::run(bool initializerKernels, Forwarding mode) {
    if (!cruncher_ || cruncher_->errorCode() != 0) return;
    Cores& cores = *cruncher_->numberCruncher();
    auto refs = parameterRefs(mode);
    if (refs.empty()) {
        fprintf(stderr, "no buffer found.\n");
        return;
    }
    const auto& list = initializerKernels ? initKernels_ : kernels_;
    const size_t n = list.size();
    for (size_t i = 0; i < n; i++) {
        std::vector<ArrayRef> r = refs;
        if (mode == Forwarding::Host && n > 1) {
            // inputs are uploaded by the first kernel only, outputs downloaded by the last one only (ClPipeline.cs:421-520)
            const size_t nin = inputs_.size(), nout = outputs_.size();
            if (i > 0)
                for (size_t j = 0; j < nin; j++) r[j].flags = 0u;
            if (i + 1 < n)
                for (size_t j = 0; j < nout; j++) r[r.size() - nout + j].flags = 0u;
        }
        // an output whose length is not the kernel's range times its elements per work-item — a reduction result such as a
        // histogram, or a frame written by a kernel whose work-items are not pixels (the strip kernels) — cannot be
        // downloaded as one slice per work-item: it comes back whole ("write all", Worker.cs:1353)
        for (auto& ar : r)
            if ((ar.flags & F_WRITE) && ar.length != (int64_t)list[i].globalRange * ar.elemsPerWorkItem) ar.flags |= F_ALL;
        const int computeId = initializerKernels ? (int)(1000000 + i) : (int)(i + 1);
        cores.compute({list[i].name}, cruncher_->repeatCount, std::string(), r, list[i].globalRange, computeId, 0, false, 4,
                      Cores::PIPELINE_EVENT, list[i].localRange);
    }
    if (initializerKernels) cruncher_->sync();
}

// ------------------------------------------------------------------ pipeline
int ClPipeline::errorCode() const {
    int e = 0;
    for (auto* s : stages_) e += s->errorCode();
    return e;
}
std::string ClPipeline::errorMessage() const {
    std::string m;
    for (auto* s : stages_) m += s->errorMessage();
    return m;
}

bool ClPipeline::pushData(const std::vector<void*>& data, const std::vector<void*>& popResultsHere) {
    const int S = (int)stages_.size();
    if (S == 0 || errorCode() != 0) {
        fprintf(stderr, "Zero pipeline stages or a stage failed to build.\n%s\n", errorMessage().c_str());
        return false;
    }
    const bool hasData = !data.empty(), hasPop = !popResultsHere.empty();
    if (chained_) return pushChained(data, popResultsHere);

    // 1) every stage computes on its `buf`s — enqueued without host synchronisation, so all devices run concurrently
    for (auto* s : stages_) {
        s->cruncher_->setEnqueueMode(true);
        s->run(false, mode_);
    }

    // 2) meanwhile, every stage's previous results move on: duplicate output(stage k) → duplicate input(stage k+1)
    for (int k = 0; k < S; k++) {
        ClPipelineStage* s = stages_[k];
        if (k == 0 && hasData) {
            for (size_t j = 0; j < s->inputs_.size() && j < data.size(); j++) {
                ClArray* dst = s->inputs_[j]->bufDuplicate();
                if (mode_ == Forwarding::PeerToPeer) {
                    ck_handle b = s->cruncher_->numberCruncher()->worker(0).buffer(dst->ref());
                    writeToBuffer(s->forwardQueue_, b, data[j]);
                } else {
                    memcpy(dst->data(), data[j], (size_t)(dst->Length() * dst->elementSize()));
                }
            }
        }
        if (k + 1 < S) {
            ClPipelineStage* nx = stages_[k + 1];
            for (size_t j = 0; j < s->outputs_.size() && j < nx->inputs_.size(); j++) {
                ClArray* src = s->outputs_[j]->bufDuplicate();
                ClArray* dst = nx->inputs_[j]->bufDuplicate();
                const int64_t n = std::min(src->Length(), dst->Length());
                if (mode_ == Forwarding::PeerToPeer) {
                    ck_handle sb = s->cruncher_->numberCruncher()->worker(0).buffer(src->ref());
                    ck_handle db = nx->cruncher_->numberCruncher()->worker(0).buffer(dst->ref());
                    ckCopyBufferPeer(nx->forwardQueue_, db, sb, 0, (int32_t)n, nullptr, nullptr);  // NVLink, no host bounce
                } else {
                    memcpy(dst->data(), src->data(), (size_t)(n * src->elementSize()));
                }
            }
        } else if (hasPop) {
            for (size_t j = 0; j < s->outputs_.size() && j < popResultsHere.size(); j++) {
                ClArray* src = s->outputs_[j]->bufDuplicate();
                if (mode_ == Forwarding::PeerToPeer) {
                    ck_handle b = s->cruncher_->numberCruncher()->worker(0).buffer(src->ref());
                    readFromBuffer(s->forwardQueue_, b, popResultsHere[j]);
                } else {
                    memcpy(popResultsHere[j], src->data(), (size_t)(src->Length() * src->elementSize()));
                }
            }
        }
    }

    // 3) join: all computes and all hand-offs of this push
    for (auto* s : stages_) {
        s->cruncher_->setEnqueueMode(false);  // drains the queues the stage used
        finish(s->forwardQueue_);
    }

    // 4) switch duplicates with real buffers (ClPipeline.cs:87-111)
    for (int i = 0; i < S; i++) {
        if (hasData || i != 0)
            for (auto& b : stages_[i]->inputs_) b->switchBuffers();
        if (hasPop || i != S - 1)
            for (auto& b : stages_[i]->outputs_) b->switchBuffers();
    }

    return finishPush(hasData, hasPop);
}

bool ClPipeline::finishPush(bool hasData, bool hasPop) {
    const int S = (int)stages_.size();
    counter++;
    if ((counter > (S * 2 - 2)) && !hasData && !hasPop) return true;
    if ((counter > (S * 2 - 1)) && hasData && !hasPop) return true;
    if ((counter > (S * 2 - 1)) && !hasData && hasPop) return true;
    if ((counter > (S * 2)) && hasData && hasPop) return true;
    return false;
}

ClPipeline::~ClPipeline() {
    if (!chained_) return;
    for (auto* s : stages_) {          // nothing of this pipeline is left in flight when its stages are reused or destroyed
        if (s->cruncher_) s->cruncher_->setEnqueueMode(false);
        if (s->forwardQueue_) finish(s->forwardQueue_);
    }
}

// Peer-to-peer forwarding without a host join per push.  What a push touches (ClPipeline.cs:49-125): stage k's kernels read
// its input `buf`s and write its output `buf`s; the hand-off k -> k+1 copies k's output DUPLICATES (= the `buf`s of the push
// before, the pointers are swapped at the end of every push) into k+1's input duplicates (= its `buf`s of the next push).
// So nothing inside one push depends on anything else in it, and push p depends on push p-1 only through
//     kernels(k, p)      after  hand-off into k (p-1)  [its input]   and  hand-off out of k (p-1)  [still reading the buffer it overwrites]
//     hand-off(k->k+1, p) after  kernels(k, p-1)       [its source]   and  kernels(k+1, p-1)        [still reading the buffer it overwrites]
// (upload into stage 0 and download from the last stage count as hand-offs).  These are event waits on the devices; the
// host waits only for what it must see itself — the upload of `data` (the caller may reuse those arrays on return) and the
// download into `popResultsHere` — and, to bound how far it runs ahead, for the kernels of the push BEFORE this one.  The
// kernels of this push are still running when pushData returns: the next push is prepared behind them.
bool ClPipeline::pushChained(const std::vector<void*>& data, const std::vector<void*>& popResultsHere) {
    const int S = (int)stages_.size();
    const bool hasData = !data.empty(), hasPop = !popResultsHere.empty();
    const int cur = counter & 1, prev = cur ^ 1;

    for (int k = 0; k < S; k++) {
        ClPipelineStage* s = stages_[k];
        ck_handle q = s->cruncher_->numberCruncher()->worker(0).mainQueue();
        ckQueueWaitEvent(q, s->inDone_[prev]);
        ckQueueWaitEvent(q, k + 1 < S ? stages_[k + 1]->inDone_[prev] : s->outDone_[prev]);
        s->cruncher_->setEnqueueMode(true);
        s->run(false, mode_);
        ckQueueRecordEvent(q, s->compDone_[cur]);
    }
    for (int k = 0; k < S; k++) {
        ClPipelineStage* s = stages_[k];
        if (k == 0 && hasData) {
            ckQueueWaitEvent(s->forwardQueue_, s->compDone_[prev]);
            for (size_t j = 0; j < s->inputs_.size() && j < data.size(); j++) {
                ck_handle b = s->cruncher_->numberCruncher()->worker(0).buffer(s->inputs_[j]->bufDuplicate()->ref());
                writeToBuffer(s->forwardQueue_, b, data[j]);
            }
            ckQueueRecordEvent(s->forwardQueue_, s->inDone_[cur]);
        }
        if (k + 1 < S) {
            ClPipelineStage* nx = stages_[k + 1];
            ckQueueWaitEvent(nx->forwardQueue_, s->compDone_[prev]);
            ckQueueWaitEvent(nx->forwardQueue_, nx->compDone_[prev]);
            for (size_t j = 0; j < s->outputs_.size() && j < nx->inputs_.size(); j++) {
                ClArray* src = s->outputs_[j]->bufDuplicate();
                ClArray* dst = nx->inputs_[j]->bufDuplicate();
                const int64_t n = std::min(src->Length(), dst->Length());
                ck_handle sb = s->cruncher_->numberCruncher()->worker(0).buffer(src->ref());
                ck_handle db = nx->cruncher_->numberCruncher()->worker(0).buffer(dst->ref());
                ckCopyBufferPeer(nx->forwardQueue_, db, sb, 0, (int32_t)n, nullptr, nullptr);  // NVLink, no host bounce
            }
            ckQueueRecordEvent(nx->forwardQueue_, nx->inDone_[cur]);
        } else if (hasPop) {
            // its own queue: the first stage's upload (same queue when S == 1) must not hold the download back
            ckQueueWaitEvent(s->forwardQueue_, s->compDone_[prev]);
            for (size_t j = 0; j < s->outputs_.size() && j < popResultsHere.size(); j++) {
                ck_handle b = s->cruncher_->numberCruncher()->worker(0).buffer(s->outputs_[j]->bufDuplicate()->ref());
                readFromBuffer(s->forwardQueue_, b, popResultsHere[j]);
            }
            ckQueueRecordEvent(s->forwardQueue_, s->outDone_[cur]);
        }
    }
    // what the host itself must see, and a bound on its lead
    if (hasData) finish(stages_[0]->forwardQueue_);
    if (hasPop) finish(stages_[S - 1]->forwardQueue_);
    for (auto* s : stages_) ckEventSynchronize(s->compDone_[prev]);

    for (int i = 0; i < S; i++) {
        if (hasData || i != 0)
            for (auto& b : stages_[i]->inputs_) b->switchBuffers();
        if (hasPop || i != S - 1)
            for (auto& b : stages_[i]->outputs_) b->switchBuffers();
    }
    return finishPush(hasData, hasPop);
}

}  // namespace cekirdekler

// ====================================================================================================================
// SingleGPUPipeline (ClPipeline.cs:2357-3240)
// ====================================================================================================================
namespace cekirdekler {

DevicePipelineArray::DevicePipelineArray(DevicePipelineArrayType t, ClArray* userArray) : type(t), user_(userArray) {}
DevicePipelineArray::DevicePipelineArray(DevicePipelineArrayType t, int arrType, int64_t n, int e)
    : type(t), a_(new ClArray(arrType, n)) {
    a_->numberOfElementsPerWorkItem = e;
    if (t == DevicePipelineArrayType::TRANSITION) {
        b_.reset(new ClArray(arrType, n));
        b_->numberOfElementsPerWorkItem = e;
    }
}

DevicePipeline::DevicePipeline(int deviceOrdinal, const std::string& kernelCodes, int computeQueueConcurrency)
    : device_(deviceOrdinal), code_(kernelCodes), concurrency_(computeQueueConcurrency) {}

void DevicePipeline::addStage(DevicePipelineStage* stage) {
    const int idx = (int)stages_.size();
    stages_.push_back(stage);
    for (auto* a : stage->arrays)
        if (a->type == DevicePipelineArrayType::TRANSITION && a->producerStage < 0) a->producerStage = idx;
}

// Every stage is one ClNumberCruncher.compute() in enqueue mode on its own queue (enqueueModeAsyncEnable advances the
// queue cursor, Cores.cs:83), carrying its own uploads (INPUT: whole `read`), kernels and downloads (OUTPUT: `write`).
// In parallel mode stage k consumes what stage k-1 produced during the previous feed(), so all stages of one feed()
// are independent and overlap on the device (ClPipeline.cs:2448-2536); in serial mode they run in order on one feed().
void DevicePipeline::enqueueAll() {
    if (!cruncher_) cruncher_.reset(new ClNumberCruncher({device_}, code_, false, concurrency_));
    if (cruncher_->errorCode() != 0 || !cruncher_->numberCruncher()) return;
    Cores& cores = *cruncher_->numberCruncher();
    cores.setEnqueueMode(true);
    for (int i = 0; i < (int)stages_.size(); i++) {
        DevicePipelineStage& st = *stages_[i];
        std::vector<ArrayRef> refs;
        for (auto* a : st.arrays) {
            int v = 0;
            if (a->type == DevicePipelineArrayType::TRANSITION && !serialMode_)
                v = (a->producerStage == i) ? (runCounter_ & 1) : ((runCounter_ + 1) & 1);
            ArrayRef r = a->version(v)->ref();
            r.flags = 0u;
            if (!st.stopHostDeviceTransmission) {
                if (a->type == DevicePipelineArrayType::INPUT) r.flags = F_READ;
                if (a->type == DevicePipelineArrayType::OUTPUT) r.flags = F_WRITE;
            }
            refs.push_back(r);
        }
        if (!serialMode_) cores.setEnqueueModeAsyncEnable(true);  // next queue for this stage
        std::vector<std::string> names;
        {
            std::string cur;
            for (char c : st.kernelNames + " ") {
                if (c == ' ' || c == ',' || c == ';' || c == '-' || c == '\n' || c == '@') {
                    if (!cur.empty()) names.push_back(cur);
                    cur.clear();
                } else {
                    cur.push_back(c);
                }
            }
        }
        cores.compute(names, 1, std::string(), refs, st.globalRange, /*computeId=*/i, 0, false, 4, Cores::PIPELINE_EVENT, st.localRange);
        if (!serialMode_) {
            cores.flushLastUsedCommandQueue();
            cores.setEnqueueModeAsyncEnable(false);
        }
    }
    open_ = true;
}

void DevicePipeline::feedAsyncBegin() { enqueueAll(); }
void DevicePipeline::feedAsyncEnd() {
    if (!open_ || !cruncher_ || !cruncher_->numberCruncher()) return;
    cruncher_->numberCruncher()->setEnqueueMode(false);  // joins every queue used by this feed
    open_ = false;
    runCounter_++;
}
void DevicePipeline::feed() {
    feedAsyncBegin();
    feedAsyncEnd();
}

}  // namespace cekirdekler
