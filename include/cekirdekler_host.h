/*
 * cekirdekler_host.h — flat C view of libcekirdekler_host.so, the C++ host layer that stands where the
 * reference's C# classes stand (there is no CLR on the build or GPU machines: SURVEY.md F6/F7).
 *
 * Each group mirrors one reference class; the file:line is the C# member it stands for (paths relative to
 * /root/reference/Cekirdekler/Cekirdekler/).  Language bindings (the Python package cekirdekler_b200, via ctypes) wrap these
 * into classes with the reference's names.  The library reaches the GPU only through include/KutuphaneCL.h.
 *
 * All functions are thread-compatible per object (the same rules as the C# classes they mirror); none throws.
 */
#ifndef CEKIRDEKLER_HOST_H
#define CEKIRDEKLER_HOST_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* ckh_cruncher;  /* ClNumberCruncher */
typedef void* ckh_array;     /* ClArray<T> */

/* ---- Hardware.ClPlatforms.all().gpus() ------------------------------ ClObjectApi.cs:204, :390-404 */
int32_t     ckh_gpu_count(void);
const char* ckh_gpu_name(int32_t ordinal);              /* borrowed, valid until the next call on this thread */
int32_t     ckh_gpu_compute_units(int32_t ordinal);
uint64_t    ckh_gpu_memory_size(int32_t ordinal);

/* ---- ClArray<T> / FastArr<T> ---------------------------------------- ClArray.cs:717-1907, CSpaceArrays.cs:234-577 */
/* arrType: CSpaceArrays.ARR_* (FLOAT 0, DOUBLE 1, LONG 2, INT 3, UINT 4, BYTE 5, CHAR 6) */
ckh_array ckh_array_create(int32_t arrType, int64_t n, int32_t alignmentBytes);   /* fastArr = true  */
ckh_array ckh_array_wrap(void* userMemory, int32_t arrType, int64_t n);           /* a C# T[] */
void      ckh_array_destroy(ckh_array a);
void*     ckh_array_data(ckh_array a);
int64_t   ckh_array_length(ckh_array a);
int32_t   ckh_array_is_fast(ckh_array a);
/* flag ids: 0 read, 1 partialRead, 2 write, 3 writeAll, 4 readOnly, 5 writeOnly, 6 zeroCopy (bools),
 *           7 numberOfElementsPerWorkItem, 8 alignmentBytes (ints)                ClArray.cs:1742-1888 */
void      ckh_array_set(ckh_array a, int32_t flagId, int32_t value);
int32_t   ckh_array_get(ckh_array a, int32_t flagId);
const char* ckh_array_flag_string(ckh_array a);          /* ClArray.cs:1685-1703; borrowed (thread-local) */

/* ---- ClNumberCruncher ------------------------------------------------ ClNumberCruncher.cs:313-349 */
ckh_cruncher ckh_cruncher_create(const int32_t* deviceOrdinals, int32_t nDevices, const char* kernelString,
                                 int32_t noPipelining, int32_t computeQueueConcurrency, int32_t streamingEnabled);
void        ckh_cruncher_destroy(ckh_cruncher c);
int32_t     ckh_cruncher_error_code(ckh_cruncher c);                 /* :380-385 */
const char* ckh_cruncher_error_message(ckh_cruncher c);              /* :387-392 */
int32_t     ckh_cruncher_errors_happened(ckh_cruncher c);
int32_t     ckh_cruncher_num_devices(ckh_cruncher c);
const char* ckh_cruncher_device_name(ckh_cruncher c, int32_t i);     /* :296-299 */
int32_t     ckh_cruncher_num_kernels(const char* kernelString);      /* kernel-name regex, :219-228 */
const char* ckh_cruncher_kernel_name(const char* kernelString, int32_t i);
/* property ids: 0 enqueueMode, 1 enqueueModeAsyncEnable, 2 noComputeMode, 3 fineGrainedQueueControl, 4 repeatCount,
 *   5 performanceFeed, 6 smoothLoadBalancer, 7 balancerMode (0 reference, 1 adaptive, 2 profile), 8 deviceClock,
 *   9 broadcastWholeReads                                            ClNumberCruncher.cs:66-187 */
void        ckh_cruncher_set(ckh_cruncher c, int32_t propId, int32_t value);
int32_t     ckh_cruncher_get(ckh_cruncher c, int32_t propId);
void        ckh_cruncher_set_repeat_kernel_name(ckh_cruncher c, const char* name);
int32_t     ckh_cruncher_markers_remaining(ckh_cruncher c);          /* :356-361 */
int32_t     ckh_cruncher_markers_reached(ckh_cruncher c);            /* :367-372 */
void        ckh_cruncher_flush_last_used_queue(ckh_cruncher c);      /* :141-145 */
void        ckh_cruncher_sync(ckh_cruncher c);
const char* ckh_cruncher_performance_report(ckh_cruncher c, int32_t computeId);   /* Cores.cs:994-1063 */
/* Device address of the buffer that device `deviceIndex` of the cruncher keeps for array `a` (created on first use, like
 * compute() does; 0 on error or for zero-copy arrays).  No reference counterpart: it feeds the peer-pointer tables of
 * kernels that write into the other devices' buffers over NVLink (nBody3Dwp). */
uint64_t    ckh_cruncher_buffer_pointer(ckh_cruncher c, ckh_array a, int32_t deviceIndex);
/* Cores.performanceHistory (Cores.cs:1075-1098): 10 rows x devices normalised shares, oldest row first, row-major into
 * `out` (capacity in doubles); returns the number of devices, 0 if `out` is too small */
int32_t     ckh_cruncher_performance_history(ckh_cruncher c, int32_t computeId, double* out, int32_t capacity);
/* fill `out` (capacity n) and return the number of devices the range is split over                  */
int32_t     ckh_cruncher_ranges(ckh_cruncher c, int32_t computeId, int32_t* out, int32_t n);            /* Cores.globalRanges */
int32_t     ckh_cruncher_references(ckh_cruncher c, int32_t computeId, int32_t* out, int32_t n);        /* Cores.globalReferences */
int32_t     ckh_cruncher_benchmarks(ckh_cruncher c, int32_t computeId, double* out, int32_t n);         /* Cores.benchmarks, ms */
int32_t     ckh_cruncher_normalized_compute_powers(ckh_cruncher c, double* out, int32_t n);             /* :254-270 */
int32_t     ckh_cruncher_normalized_global_ranges(ckh_cruncher c, int32_t computeId, double* out, int32_t n); /* :277-294 */
void        ckh_cruncher_reset_balancer(ckh_cruncher c, int32_t computeId);
/* SPMD (one process per GPU): this cruncher's single device is global device `rank` of `world`;
 * `exchange(times, world, user)` must all-gather the per-rank device times in place. */
typedef void (*ckh_exchange_fn)(double* times, int32_t world, void* user);
void        ckh_cruncher_set_spmd(ckh_cruncher c, int32_t rank, int32_t world, ckh_exchange_fn exchange, void* user);

/* ---- ClParameterGroup.compute / ClArray.compute ---------------------- ClArray.cs:543-659, :1605-1736 */
/* readWritesReady / elementsPerItemReady may be NULL (then each array's own flags are used).  Returns 0 when the
 * compute was dispatched, otherwise the number of the validation rule that refused it. */
int32_t ckh_compute(ckh_cruncher c, ckh_array* arrays, int32_t nArrays, int32_t computeId, const char* kernelNamesString,
                    int32_t globalRange, int32_t localRange, int32_t ofsetGlobalRange, int32_t pipeline, int32_t pipelineType,
                    int32_t pipelineBlobs, const char** readWritesReady, const int32_t* elementsPerItemReady);

/* ---- Functions.loadBalance (pure host arithmetic) -------------------- HelperFunctions.cs:190-280, Cores.cs:569-605 */
typedef void* ckh_balancer;
ckh_balancer ckh_balancer_create(int32_t devices, int32_t mode /* 0 reference, 1 adaptive, 2 profile */);
void         ckh_balancer_destroy(ckh_balancer b);
/* One Cores.compute-style step: on the first call (all ranges 0) does the equal split + bench=10 pass; later calls use
 * `benchmarkMs`.  `ranges` (in/out, `devices` ints) sums to totalRange afterwards. */
void         ckh_balancer_step(ckh_balancer b, const double* benchmarkMs, int32_t smooth, int32_t totalRange, int32_t* ranges,
                               int32_t pipelineEnabled, int32_t blobs, int32_t localRange);
/* the raw reference function on caller-held state: history is kHistoryDepth(10) x devices doubles, oldest row first */
void         ckh_load_balance(const double* benchmarkMs, int32_t smooth, double* history, int32_t totalRange, int32_t* ranges,
                              int32_t step, int32_t devices);

/* ---- Pipeline.ClPipeline / ClPipelineStage ---------------------------- ClPipeline.cs:41-2356 */
typedef void* ckh_stage;     /* ClPipelineStage */
typedef void* ckh_pipeline;  /* ClPipeline */
ckh_stage    ckh_stage_create(int32_t debugLog);                                                         /* :176 */
void         ckh_stage_destroy(ckh_stage s);
void         ckh_stage_add_devices(ckh_stage s, const int32_t* deviceOrdinals, int32_t n);               /* :1731 */
void         ckh_stage_add_kernels(ckh_stage s, const char* kernels, const char* kernelNames, const int32_t* globalRanges,
                                   const int32_t* localRanges, int32_t n);                               /* :1744 */
void         ckh_stage_initializer_kernel(ckh_stage s, const char* kernelNames, const int32_t* globalRanges,
                                          const int32_t* localRanges, int32_t n);                        /* :1678 */
/* role: 0 input (:1810), 1 output (:1845), 2 hidden (:1777); the stage allocates `buf` and `bufDuplicate` */
void         ckh_stage_add_buffer(ckh_stage s, int32_t role, int32_t arrType, int64_t n, int32_t elementsPerWorkItem);
void         ckh_stage_prepend_to(ckh_stage s, ckh_stage other);   /* s runs before other  (:1704) */
void         ckh_stage_append_to(ckh_stage s, ckh_stage other);    /* s runs after other   (:1717) */
void         ckh_stage_set_enqueue_mode(ckh_stage s, int32_t v);
/* hostForwarding = 0: hand-offs are NVLink peer copies between device buffers; 1: the reference's host bounce (:1422-1574) */
ckh_pipeline ckh_stage_make_pipeline(ckh_stage s, int32_t hostForwarding);                               /* :1582 */
void         ckh_pipeline_destroy(ckh_pipeline p);
/* pushData(data, popResultsHere) (:49-125): host arrays or NULL/0; returns 1 once the delivered results are valid */
int32_t      ckh_pipeline_push(ckh_pipeline p, void** data, int32_t nData, void** popResultsHere, int32_t nResults);
int32_t      ckh_pipeline_error_code(ckh_pipeline p);
const char*  ckh_pipeline_error_message(ckh_pipeline p);
int32_t      ckh_pipeline_forwarding(ckh_pipeline p);   /* 0 peer-to-peer, 1 host */
int32_t      ckh_pipeline_counter(ckh_pipeline p);

/* ---- Pipeline.SingleGPUPipeline: stages overlapped on one device ------ ClPipeline.cs:2357-3240 */
typedef void* ckh_dparray;    /* DevicePipelineArray */
typedef void* ckh_dpstage;    /* DevicePipelineStage */
typedef void* ckh_dpipeline;  /* DevicePipeline */
/* type: DevicePipelineArrayType INPUT 0, OUTPUT 1, INTERNAL 2, TRANSITION 3 (:3178-3184) */
ckh_dparray   ckh_dparray_wrap(int32_t type, ckh_array userArray);                                  /* INPUT / OUTPUT over a caller array */
ckh_dparray   ckh_dparray_create(int32_t type, int32_t arrType, int64_t n, int32_t elementsPerWorkItem);  /* INTERNAL / TRANSITION */
void          ckh_dparray_destroy(ckh_dparray a);
ckh_dpstage   ckh_dpstage_create(const char* kernelNames, int32_t globalRange, int32_t localRange);  /* :3098 */
void          ckh_dpstage_destroy(ckh_dpstage s);
void          ckh_dpstage_bind_array(ckh_dpstage s, ckh_dparray a);                                  /* :3106 */
void          ckh_dpstage_stop_transmission(ckh_dpstage s, int32_t v);                               /* :2638 */
ckh_dpipeline ckh_dpipeline_create(int32_t deviceOrdinal, const char* kernelCodes, int32_t computeQueueConcurrency); /* :2369 */
void          ckh_dpipeline_destroy(ckh_dpipeline p);
void          ckh_dpipeline_add_stage(ckh_dpipeline p, ckh_dpstage s);                               /* :2381 */
void          ckh_dpipeline_feed(ckh_dpipeline p);                                                   /* :2570 */
void          ckh_dpipeline_feed_async_begin(ckh_dpipeline p);                                       /* :2603 */
void          ckh_dpipeline_feed_async_end(ckh_dpipeline p);                                         /* :2616 */
void          ckh_dpipeline_set_serial(ckh_dpipeline p, int32_t serial);                             /* :2621-2628 */
int32_t       ckh_dpipeline_error_code(ckh_dpipeline p);
const char*   ckh_dpipeline_error_message(ckh_dpipeline p);
int32_t       ckh_dpipeline_latency(ckh_dpipeline p);   /* feeds between an input and its result: 0 serial, stages-1 parallel */

/* ---- Pipeline.Pool: ClTask / ClTaskPool / ClDevicePool ----------------- ClPipeline.cs:3241-5078 */
typedef void* ckh_task;        /* ClTask */
typedef void* ckh_taskpool;    /* ClTaskPool */
typedef void* ckh_devicepool;  /* ClDevicePool */
typedef void (*ckh_callback_fn)(void* user);
/* ClArray.task()/ClParameterGroup.task() (ClArray.cs:515-521, :1552-1583): freezes a compute() call; readWrite/elems may be
 * NULL (then the arrays' current flags are frozen) */
ckh_task     ckh_task_create(ckh_array* arrays, int32_t nArrays, const char** readWrite, const int32_t* elementsPerItem,
                             int32_t computeId, const char* kernelNamesString, int32_t globalRange, int32_t localRange,
                             int32_t ofsetGlobalRange, int32_t pipeline, int32_t pipelineType, int32_t pipelineBlobs);
void         ckh_task_destroy(ckh_task t);
void         ckh_task_set_type(ckh_task t, int32_t clTaskTypeBits);      /* ClTaskType bit flags, :3247-3321 */
int32_t      ckh_task_get_type(ckh_task t);
void         ckh_task_set_callback(ckh_task t, ckh_callback_fn fn, void* user);   /* :3481 */
int32_t      ckh_task_compute(ckh_task t, ckh_cruncher c);               /* :3359 */
ckh_taskpool ckh_taskpool_create(void);
void         ckh_taskpool_destroy(ckh_taskpool p);
void         ckh_taskpool_feed(ckh_taskpool p, ckh_task t);              /* :3649 (the task is copied) */
void         ckh_taskpool_reset(ckh_taskpool p);                         /* :3642 */
int32_t      ckh_taskpool_remaining(ckh_taskpool p);                     /* :3758 */
/* poolType: 0 DEVICE_COMPUTE_AT_WILL, 1 DEVICE_ROUND_ROBIN (:3802-3808) */
ckh_devicepool ckh_devicepool_create(int32_t poolType, const char* kernelCodeToCompile, int32_t fineGrainedQueueControl,
                                     int32_t maxQueuesPerDevice);        /* :3935 */
void         ckh_devicepool_destroy(ckh_devicepool p);
void         ckh_devicepool_add_device(ckh_devicepool p, int32_t deviceOrdinal);   /* :4340 */
void         ckh_devicepool_enqueue(ckh_devicepool p, ckh_taskpool pool);          /* :4400 */
double       ckh_devicepool_finish(ckh_devicepool p);                    /* :4470, elapsed ms */
void         ckh_devicepool_dispose(ckh_devicepool p);                   /* :4527 */
int32_t      ckh_devicepool_error_code(ckh_devicepool p);
const char*  ckh_devicepool_error_message(ckh_devicepool p);
int32_t      ckh_devicepool_tasks_per_device(ckh_devicepool p, int32_t* out, int32_t n);

#ifdef __cplusplus
}
#endif
#endif /* CEKIRDEKLER_HOST_H */
