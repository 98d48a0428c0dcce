// capi.cpp — flat C view of the host layer (include/cekirdekler_host.h) for language bindings.
#include <cekirdekler_host.h>

#include "cruncher.h"

#include <cstring>

using namespace cekirdekler;

namespace {
thread_local std::string t_str;
const char* keep(const std::string& s) {
    t_str = s;
    return t_str.c_str();
}
struct BalancerBox {
    BalanceState st;
    BalancerMode mode;
    int devices;
};
template <typename T, typename V> int32_t copy_out(const V& v, T* out, int32_t n) {
    for (int32_t i = 0; i < n && i < (int32_t)v.size(); i++) out[i] = (T)v[i];
    return (int32_t)v.size();
}
}  // namespace

extern "C" {

// ------------------------------------------------------------------ devices
int32_t ckh_gpu_count(void) { return ckDeviceCount(); }
static ck_handle open_device(int32_t ordinal, ck_handle* plat) {
    ck_handle list = platformList();
    *plat = createPlatform(list, 0);
    deletePlatformList(list);
    return createDevice(*plat, CODE_GPU(), ordinal);
}
const char* ckh_gpu_name(int32_t ordinal) {
    ck_handle plat, dev = open_device(ordinal, &plat);
    std::string name;
    if (dev) {
        ck_handle s = createString();
        getDeviceName(dev, s);
        name = readFromString(s);
        deleteString(s);
        deleteDevice(dev);
    }
    deletePlatform(plat);
    return keep(name);
}
int32_t ckh_gpu_compute_units(int32_t ordinal) {
    ck_handle plat, dev = open_device(ordinal, &plat);
    int32_t r = dev ? deviceComputeUnits(dev) : 0;
    if (dev) deleteDevice(dev);
    deletePlatform(plat);
    return r;
}
uint64_t ckh_gpu_memory_size(int32_t ordinal) {
    ck_handle plat, dev = open_device(ordinal, &plat);
    uint64_t r = dev ? deviceMemSize(dev) : 0;
    if (dev) deleteDevice(dev);
    deletePlatform(plat);
    return r;
}

// ------------------------------------------------------------------ arrays
ckh_array ckh_array_create(int32_t arrType, int64_t n, int32_t alignmentBytes) {
    if (n < 0 || n > 0x7fffffffLL || arrTypeSize(arrType) == 0) return nullptr;
    auto* a = new ClArray(arrType, n, alignmentBytes > 0 ? alignmentBytes : 4096);
    if (!a->data()) {
        delete a;
        return nullptr;
    }
    return a;
}
ckh_array ckh_array_wrap(void* userMemory, int32_t arrType, int64_t n) {
    if (!userMemory || n < 0 || n > 0x7fffffffLL || arrTypeSize(arrType) == 0) return nullptr;
    return new ClArray(userMemory, arrType, n);
}
void ckh_array_destroy(ckh_array a) { delete static_cast<ClArray*>(a); }
void* ckh_array_data(ckh_array a) { return a ? static_cast<ClArray*>(a)->data() : nullptr; }
int64_t ckh_array_length(ckh_array a) { return a ? static_cast<ClArray*>(a)->Length() : 0; }
int32_t ckh_array_is_fast(ckh_array a) { return a && static_cast<ClArray*>(a)->isFastArr(); }
void ckh_array_set(ckh_array h, int32_t id, int32_t v) {
    if (!h) return;
    ClArray& a = *static_cast<ClArray*>(h);
    switch (id) {
        case 0: a.setRead(v != 0); break;
        case 1: a.setPartialRead(v != 0); break;
        case 2: a.setWrite(v != 0); break;
        case 3: a.setWriteAll(v != 0); break;
        case 4: a.setReadOnly(v != 0); break;
        case 5: a.setWriteOnly(v != 0); break;
        case 6: a.zeroCopy = (v != 0); break;
        case 7: a.numberOfElementsPerWorkItem = v; break;
        case 8: a.alignmentBytes = v; break;
    }
}
int32_t ckh_array_get(ckh_array h, int32_t id) {
    if (!h) return 0;
    ClArray& a = *static_cast<ClArray*>(h);
    switch (id) {
        case 0: return a.read();
        case 1: return a.partialRead();
        case 2: return a.write();
        case 3: return a.writeAll();
        case 4: return a.readOnly();
        case 5: return a.writeOnly();
        case 6: return a.zeroCopy;
        case 7: return a.numberOfElementsPerWorkItem;
        case 8: return a.alignmentBytes;
    }
    return 0;
}
const char* ckh_array_flag_string(ckh_array a) { return keep(a ? static_cast<ClArray*>(a)->flagString() : ""); }

// ------------------------------------------------------------------ cruncher
ckh_cruncher ckh_cruncher_create(const int32_t* ords, int32_t n, const char* src, int32_t noPipelining, int32_t concurrency,
                                 int32_t streaming) {
    std::vector<int> d(ords, ords + (n > 0 ? n : 0));
    return new ClNumberCruncher(d, src ? src : "", noPipelining != 0, concurrency > 0 ? concurrency : 16, streaming != 0);
}
void ckh_cruncher_destroy(ckh_cruncher c) { delete static_cast<ClNumberCruncher*>(c); }
#define CR(c) (*static_cast<ClNumberCruncher*>(c))
int32_t ckh_cruncher_error_code(ckh_cruncher c) { return c ? CR(c).errorCode() : 1; }
const char* ckh_cruncher_error_message(ckh_cruncher c) { return keep(c ? CR(c).errorMessage() : "null cruncher"); }
int32_t ckh_cruncher_errors_happened(ckh_cruncher c) { return c ? CR(c).numberOfErrorsHappened : 1; }
int32_t ckh_cruncher_num_devices(ckh_cruncher c) { return c ? CR(c).numberOfDevices() : 0; }
const char* ckh_cruncher_device_name(ckh_cruncher c, int32_t i) {
    if (!c || !CR(c).numberCruncher() || i < 0 || i >= CR(c).numberOfDevices()) return keep("");
    return keep(CR(c).numberCruncher()->worker(i).deviceName());
}
int32_t ckh_cruncher_num_kernels(const char* s) { return (int32_t)ClNumberCruncher::extractKernelNames(s ? s : "").size(); }
const char* ckh_cruncher_kernel_name(const char* s, int32_t i) {
    auto n = ClNumberCruncher::extractKernelNames(s ? s : "");
    return keep((i >= 0 && i < (int)n.size()) ? n[i] : "");
}
void ckh_cruncher_set(ckh_cruncher c, int32_t id, int32_t v) {
    if (!c) return;
    ClNumberCruncher& cr = CR(c);
    Cores* k = cr.numberCruncher();
    switch (id) {
        case 0: cr.setEnqueueMode(v != 0); break;
        case 1: cr.setEnqueueModeAsyncEnable(v != 0); break;
        case 2: cr.setNoComputeMode(v != 0); break;
        case 3: cr.setFineGrainedQueueControl(v != 0); break;
        case 4: cr.repeatCount = v; break;
        case 5: cr.performanceFeed = (v != 0); break;
        case 6: cr.setSmoothLoadBalancer(v != 0); break;
        case 7: if (k) k->balancerMode = v == 2 ? BalancerMode::Profile : (v ? BalancerMode::Adaptive : BalancerMode::Reference); break;
        case 8: if (k) k->deviceClock = (v != 0); break;
        case 9: if (k) k->broadcastWholeReads = (v != 0); break;
    }
}
int32_t ckh_cruncher_get(ckh_cruncher c, int32_t id) {
    if (!c) return 0;
    ClNumberCruncher& cr = CR(c);
    Cores* k = cr.numberCruncher();
    switch (id) {
        case 0: return cr.enqueueMode();
        case 1: return k && k->enqueueModeAsyncEnable();
        case 2: return k && k->noComputeMode;
        case 3: return k && k->fineGrainedQueueControl;
        case 4: return cr.repeatCount;
        case 5: return cr.performanceFeed;
        case 6: return k && k->smoothLoadBalancer;
        case 7: return k ? (int)k->balancerMode : 0;
        case 8: return k && k->deviceClock;
        case 9: return k && k->broadcastWholeReads;
    }
    return 0;
}
void ckh_cruncher_set_repeat_kernel_name(ckh_cruncher c, const char* name) {
    if (c) CR(c).repeatKernelName = name ? name : "";
}
int32_t ckh_cruncher_markers_remaining(ckh_cruncher c) { return c ? CR(c).countMarkersRemaining() : 0; }
int32_t ckh_cruncher_markers_reached(ckh_cruncher c) { return c ? CR(c).countMarkersReached() : 0; }
void ckh_cruncher_flush_last_used_queue(ckh_cruncher c) { if (c) CR(c).flushLastUsedCommandQueue(); }
void ckh_cruncher_sync(ckh_cruncher c) { if (c) CR(c).sync(); }
const char* ckh_cruncher_performance_report(ckh_cruncher c, int32_t id) {
    if (!c || !CR(c).numberCruncher()) return keep("Error: Number cruncher core object was not allocated properly.");
    return keep(CR(c).numberCruncher()->performanceReport(id));
}
uint64_t ckh_cruncher_buffer_pointer(ckh_cruncher c, ckh_array a, int32_t deviceIndex) {
    // device address of the buffer that device `deviceIndex` of the cruncher keeps for `a` (created on first use, as
    // compute() would); 0 on error.  Feeds the peer tables of kernels that store into other devices' buffers.
    if (!c || !a || !CR(c).numberCruncher()) return 0;
    Cores* cores = CR(c).numberCruncher();
    if (deviceIndex < 0 || deviceIndex >= cores->numberOfDevices()) return 0;
    Worker& w = cores->worker(deviceIndex);
    std::lock_guard<std::mutex> lk(w.mu);
    ck_handle b = w.buffer(static_cast<ClArray*>(a)->ref());
    if (!b || ckBufferIsZeroCopy(b)) return 0;
    return (uint64_t)(uintptr_t)ckBufferDevicePointer(b);
}
int32_t ckh_cruncher_performance_history(ckh_cruncher c, int32_t id, double* out, int32_t cap) {
    // Cores.performanceHistory (Cores.cs:1075-1098): kHistoryDepth rows of normalised throughput shares, oldest first;
    // all zero for a compute id that has not run (the reference creates the zero matrix on first access)
    if (!c || !CR(c).numberCruncher() || !out) return 0;
    auto* cores = CR(c).numberCruncher();
    auto* st = cores->state(id);
    const int devices = st && !st->balance.history.empty() ? (int)st->balance.history[0].size()
                                                            : (cores->spmdWorld() > 0 ? cores->spmdWorld() : cores->numberOfDevices());
    if (cap < kHistoryDepth * devices) return 0;
    for (int k = 0; k < kHistoryDepth; k++)
        for (int j = 0; j < devices; j++)
            out[k * devices + j] = (st && (int)st->balance.history.size() == kHistoryDepth) ? st->balance.history[k][j] : 0.0;
    return devices;
}
int32_t ckh_cruncher_ranges(ckh_cruncher c, int32_t id, int32_t* out, int32_t n) {
    if (!c || !CR(c).numberCruncher()) return 0;
    auto* st = CR(c).numberCruncher()->state(id);
    return st ? copy_out(st->ranges, out, n) : 0;
}
int32_t ckh_cruncher_references(ckh_cruncher c, int32_t id, int32_t* out, int32_t n) {
    if (!c || !CR(c).numberCruncher()) return 0;
    auto* st = CR(c).numberCruncher()->state(id);
    return st ? copy_out(st->references, out, n) : 0;
}
int32_t ckh_cruncher_benchmarks(ckh_cruncher c, int32_t id, double* out, int32_t n) {
    if (!c || !CR(c).numberCruncher()) return 0;
    return copy_out(CR(c).numberCruncher()->benchmarks(id), out, n);
}
int32_t ckh_cruncher_normalized_compute_powers(ckh_cruncher c, double* out, int32_t n) {
    if (!c || !CR(c).numberCruncher()) return 0;
    return copy_out(CR(c).numberCruncher()->normalizedComputePowers(), out, n);
}
int32_t ckh_cruncher_normalized_global_ranges(ckh_cruncher c, int32_t id, double* out, int32_t n) {
    if (!c || !CR(c).numberCruncher()) return 0;
    return copy_out(CR(c).numberCruncher()->normalizedGlobalRanges(id), out, n);
}
void ckh_cruncher_reset_balancer(ckh_cruncher c, int32_t id) {
    if (c && CR(c).numberCruncher()) CR(c).numberCruncher()->resetBalancer(id);
}
void ckh_cruncher_set_spmd(ckh_cruncher c, int32_t rank, int32_t world, ckh_exchange_fn exchange, void* user) {
    if (!c || !CR(c).numberCruncher()) return;
    Cores* k = CR(c).numberCruncher();
    k->setSpmd(rank, world);
    if (exchange) k->exchangeBenchmarks = [exchange, user](std::vector<double>& t) { exchange(t.data(), (int32_t)t.size(), user); };
    else k->exchangeBenchmarks = nullptr;
}

// ------------------------------------------------------------------ compute
int32_t ckh_compute(ckh_cruncher c, ckh_array* arrays, int32_t nArrays, int32_t computeId, const char* kernels, int32_t globalRange,
                    int32_t localRange, int32_t offset, int32_t pipeline, int32_t pipelineType, int32_t blobs, const char** rw,
                    const int32_t* elems) {
    if (!c || !arrays || nArrays <= 0) return -1;
    ClParameterGroup g;
    for (int i = 0; i < nArrays; i++) {
        if (!arrays[i]) return -1;
        g.nextParam(static_cast<ClArray*>(arrays[i]));
    }
    std::vector<std::string> rws;
    std::vector<int> eps;
    if (rw)
        for (int i = 0; i < nArrays; i++) rws.push_back(rw[i] ? rw[i] : "");
    if (elems) eps.assign(elems, elems + nArrays);
    return g.compute(CR(c), computeId, kernels ? kernels : "", globalRange, localRange, offset, pipeline != 0, pipelineType != 0, blobs,
                     rw ? &rws : nullptr, elems ? &eps : nullptr);
}

// ------------------------------------------------------------------ balancer
ckh_balancer ckh_balancer_create(int32_t devices, int32_t mode) {
    auto* b = new BalancerBox();
    b->devices = devices;
    b->mode = mode == 2 ? BalancerMode::Profile : (mode ? BalancerMode::Adaptive : BalancerMode::Reference);
    b->st.reset(devices);
    return b;
}
void ckh_balancer_destroy(ckh_balancer b) { delete static_cast<BalancerBox*>(b); }
void ckh_balancer_step(ckh_balancer h, const double* benchMs, int32_t smooth, int32_t totalRange, int32_t* ranges, int32_t pipelineEnabled,
                       int32_t blobs, int32_t localRange) {
    if (!h) return;
    auto& b = *static_cast<BalancerBox*>(h);
    std::vector<int> r(ranges, ranges + b.devices);
    bool first = true;
    for (int v : r) first = first && (v == 0);
    std::vector<double> bench(b.devices, 10.0);
    if (first) initialSplit(totalRange, b.devices, r);
    else bench.assign(benchMs, benchMs + b.devices);
    int step = balanceStep(r, pipelineEnabled != 0, blobs, localRange, totalRange);
    if (totalRange % step != 0) step = localRange;
    loadBalance(bench, smooth != 0, b.st, totalRange, r, step, b.mode);
    for (int i = 0; i < b.devices; i++) ranges[i] = r[i];
}
void ckh_load_balance(const double* benchMs, int32_t smooth, double* history, int32_t totalRange, int32_t* ranges, int32_t step,
                      int32_t devices) {
    BalanceState st;
    st.reset(devices);
    for (int k = 0; k < kHistoryDepth; k++)
        for (int j = 0; j < devices; j++) st.history[k][j] = history[k * devices + j];
    std::vector<double> bench(benchMs, benchMs + devices);
    std::vector<int> r(ranges, ranges + devices);
    loadBalance(bench, smooth != 0, st, totalRange, r, step, BalancerMode::Reference);
    for (int k = 0; k < kHistoryDepth; k++)
        for (int j = 0; j < devices; j++) history[k * devices + j] = st.history[k][j];
    for (int i = 0; i < devices; i++) ranges[i] = r[i];
}

}  // extern "C"

// ------------------------------------------------------------------ device-to-device pipeline (ClPipeline.cs:41-2356)
#include "pipeline.h"
#include "pool.h"

extern "C" {

ckh_stage ckh_stage_create(int32_t debugLog) { return new ClPipelineStage(debugLog != 0); }
void ckh_stage_destroy(ckh_stage s) { delete static_cast<ClPipelineStage*>(s); }
#define ST(s) (*static_cast<ClPipelineStage*>(s))
void ckh_stage_add_devices(ckh_stage s, const int32_t* ords, int32_t n) {
    if (s) ST(s).addDevices(std::vector<int>(ords, ords + (n > 0 ? n : 0)));
}
void ckh_stage_add_kernels(ckh_stage s, const char* kernels, const char* names, const int32_t* globals, const int32_t* locals, int32_t n) {
    if (s) ST(s).addKernels(kernels ? kernels : "", names ? names : "", std::vector<int>(globals, globals + n), std::vector<int>(locals, locals + n));
}
void ckh_stage_initializer_kernel(ckh_stage s, const char* names, const int32_t* globals, const int32_t* locals, int32_t n) {
    if (s) ST(s).initializerKernel(names ? names : "", std::vector<int>(globals, globals + n), std::vector<int>(locals, locals + n));
}
void ckh_stage_add_buffer(ckh_stage s, int32_t role, int32_t arrType, int64_t n, int32_t elementsPerWorkItem) {
    if (!s) return;
    if (role == 0) ST(s).addInputBuffer(arrType, n, elementsPerWorkItem);
    else if (role == 1) ST(s).addOutputBuffer(arrType, n, elementsPerWorkItem);
    else ST(s).addHiddenBuffer(arrType, n, elementsPerWorkItem);
}
void ckh_stage_prepend_to(ckh_stage s, ckh_stage other) { if (s) ST(s).prependToStage(static_cast<ClPipelineStage*>(other)); }
void ckh_stage_append_to(ckh_stage s, ckh_stage other) { if (s) ST(s).appendToStage(static_cast<ClPipelineStage*>(other)); }
void ckh_stage_set_enqueue_mode(ckh_stage s, int32_t v) { if (s) ST(s).enqueueMode = (v != 0); }
ckh_pipeline ckh_stage_make_pipeline(ckh_stage s, int32_t hostForwarding) {
    return s ? ST(s).makePipeline(hostForwarding ? Forwarding::Host : Forwarding::PeerToPeer) : nullptr;
}
void ckh_pipeline_destroy(ckh_pipeline p) { delete static_cast<ClPipeline*>(p); }
int32_t ckh_pipeline_push(ckh_pipeline p, void** data, int32_t nData, void** results, int32_t nResults) {
    if (!p) return 0;
    std::vector<void*> d(data, data + (data ? nData : 0)), r(results, results + (results ? nResults : 0));
    return static_cast<ClPipeline*>(p)->pushData(d, r) ? 1 : 0;
}
int32_t ckh_pipeline_error_code(ckh_pipeline p) { return p ? static_cast<ClPipeline*>(p)->errorCode() : 1; }
const char* ckh_pipeline_error_message(ckh_pipeline p) { return keep(p ? static_cast<ClPipeline*>(p)->errorMessage() : "null pipeline"); }
int32_t ckh_pipeline_forwarding(ckh_pipeline p) { return p ? (int32_t)static_cast<ClPipeline*>(p)->forwarding() : -1; }
int32_t ckh_pipeline_counter(ckh_pipeline p) { return p ? static_cast<ClPipeline*>(p)->counter : 0; }

// ---- SingleGPUPipeline (ClPipeline.cs:2357-3240)
ckh_dparray ckh_dparray_wrap(int32_t type, ckh_array userArray) {
    if (!userArray) return nullptr;
    return new DevicePipelineArray((DevicePipelineArrayType)type, static_cast<ClArray*>(userArray));
}
ckh_dparray ckh_dparray_create(int32_t type, int32_t arrType, int64_t n, int32_t elementsPerWorkItem) {
    return new DevicePipelineArray((DevicePipelineArrayType)type, arrType, n, elementsPerWorkItem);
}
void ckh_dparray_destroy(ckh_dparray a) { delete static_cast<DevicePipelineArray*>(a); }
ckh_dpstage ckh_dpstage_create(const char* kernelNames, int32_t globalRange, int32_t localRange) {
    return new DevicePipelineStage(kernelNames ? kernelNames : "", globalRange, localRange);
}
void ckh_dpstage_destroy(ckh_dpstage s) { delete static_cast<DevicePipelineStage*>(s); }
void ckh_dpstage_bind_array(ckh_dpstage s, ckh_dparray a) {
    if (s && a) static_cast<DevicePipelineStage*>(s)->bindArray(static_cast<DevicePipelineArray*>(a));
}
void ckh_dpstage_stop_transmission(ckh_dpstage s, int32_t v) {
    if (s) static_cast<DevicePipelineStage*>(s)->stopHostDeviceTransmission = (v != 0);
}
ckh_dpipeline ckh_dpipeline_create(int32_t deviceOrdinal, const char* kernelCodes, int32_t computeQueueConcurrency) {
    return new DevicePipeline(deviceOrdinal, kernelCodes ? kernelCodes : "", computeQueueConcurrency > 0 ? computeQueueConcurrency : 16);
}
void ckh_dpipeline_destroy(ckh_dpipeline p) { delete static_cast<DevicePipeline*>(p); }
#define DP(p) (*static_cast<DevicePipeline*>(p))
void ckh_dpipeline_add_stage(ckh_dpipeline p, ckh_dpstage s) { if (p && s) DP(p).addStage(static_cast<DevicePipelineStage*>(s)); }
void ckh_dpipeline_feed(ckh_dpipeline p) { if (p) DP(p).feed(); }
void ckh_dpipeline_feed_async_begin(ckh_dpipeline p) { if (p) DP(p).feedAsyncBegin(); }
void ckh_dpipeline_feed_async_end(ckh_dpipeline p) { if (p) DP(p).feedAsyncEnd(); }
void ckh_dpipeline_set_serial(ckh_dpipeline p, int32_t serial) {
    if (!p) return;
    if (serial) DP(p).enableSerialMode();
    else DP(p).enableParallelMode();
}
int32_t ckh_dpipeline_error_code(ckh_dpipeline p) { return p ? DP(p).errorCode() : 1; }
const char* ckh_dpipeline_error_message(ckh_dpipeline p) { return keep(p ? DP(p).errorMessage() : "null pipeline"); }
int32_t ckh_dpipeline_latency(ckh_dpipeline p) { return p ? DP(p).latency() : 0; }

// ---- Pool: ClTask / ClTaskPool / ClDevicePool (ClPipeline.cs:3241-5078)
ckh_task ckh_task_create(ckh_array* arrays, int32_t nArrays, const char** readWrite, const int32_t* elems, int32_t computeId,
                         const char* kernelNames, int32_t globalRange, int32_t localRange, int32_t offset, int32_t pipeline,
                         int32_t pipelineType, int32_t blobs) {
    auto* t = new ClTask();
    for (int i = 0; i < nArrays; i++) {
        ClArray* a = static_cast<ClArray*>(arrays[i]);
        t->arrays.push_back(a);
        // ClArray.task() freezes the flags at creation time (ClArray.cs:1552-1583)
        t->readWrite.push_back(readWrite && readWrite[i] ? std::string(readWrite[i]) : a->flagString());
        t->elementsPerItem.push_back(elems ? elems[i] : a->numberOfElementsPerWorkItem);
    }
    t->computeId = computeId;
    t->kernelNamesString = kernelNames ? kernelNames : "";
    t->globalRange = globalRange;
    t->localRange = localRange;
    t->ofsetGlobalRange = offset;
    t->pipeline = pipeline != 0;
    t->pipelineType = pipelineType != 0;
    t->pipelineBlobs = blobs;
    return t;
}
void ckh_task_destroy(ckh_task t) { delete static_cast<ClTask*>(t); }
void ckh_task_set_type(ckh_task t, int32_t type) { if (t) static_cast<ClTask*>(t)->type = type; }
int32_t ckh_task_get_type(ckh_task t) { return t ? static_cast<ClTask*>(t)->type : 0; }
void ckh_task_set_callback(ckh_task t, ckh_callback_fn fn, void* user) {
    if (!t) return;
    if (fn) static_cast<ClTask*>(t)->callback = [fn, user] { fn(user); };
    else static_cast<ClTask*>(t)->callback = nullptr;
}
int32_t ckh_task_compute(ckh_task t, ckh_cruncher c) { return (t && c) ? static_cast<ClTask*>(t)->compute(CR(c)) : -1; }

ckh_taskpool ckh_taskpool_create(void) { return new ClTaskPool(); }
void ckh_taskpool_destroy(ckh_taskpool p) { delete static_cast<ClTaskPool*>(p); }
void ckh_taskpool_feed(ckh_taskpool p, ckh_task t) { if (p && t) static_cast<ClTaskPool*>(p)->feed(*static_cast<ClTask*>(t)); }
void ckh_taskpool_reset(ckh_taskpool p) { if (p) static_cast<ClTaskPool*>(p)->reset(); }
int32_t ckh_taskpool_remaining(ckh_taskpool p) { return p ? static_cast<ClTaskPool*>(p)->remainingTaskGroupsOrTasks() : 0; }

ckh_devicepool ckh_devicepool_create(int32_t poolType, const char* kernelCode, int32_t fineGrained, int32_t maxQueuesPerDevice) {
    return new ClDevicePool(poolType ? DEVICE_ROUND_ROBIN : DEVICE_COMPUTE_AT_WILL, kernelCode ? kernelCode : "", fineGrained != 0,
                            maxQueuesPerDevice > 0 ? maxQueuesPerDevice : 3);
}
void ckh_devicepool_destroy(ckh_devicepool p) { delete static_cast<ClDevicePool*>(p); }
#define POOL(p) (*static_cast<ClDevicePool*>(p))
void ckh_devicepool_add_device(ckh_devicepool p, int32_t ordinal) { if (p) POOL(p).addDevice(ordinal); }
void ckh_devicepool_enqueue(ckh_devicepool p, ckh_taskpool tp) { if (p && tp) POOL(p).enqueueTaskPool(*static_cast<ClTaskPool*>(tp)); }
double ckh_devicepool_finish(ckh_devicepool p) { return p ? POOL(p).finish() : 0.0; }
void ckh_devicepool_dispose(ckh_devicepool p) { if (p) POOL(p).dispose(); }
int32_t ckh_devicepool_error_code(ckh_devicepool p) { return p ? POOL(p).errorCode() : 1; }
const char* ckh_devicepool_error_message(ckh_devicepool p) { return keep(p ? POOL(p).errorMessage() : "null pool"); }
int32_t ckh_devicepool_tasks_per_device(ckh_devicepool p, int32_t* out, int32_t n) {
    if (!p) return 0;
    return copy_out(POOL(p).tasksPerDevice(), out, n);
}

}  // extern "C"
