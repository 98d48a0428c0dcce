/*
 * KutuphaneCL.h — the drop-in C ABI of libKutuphaneCL.so (CUDA, sm_100a).
 *
 * This header declares the 81 cdecl symbols that the reference's C# layer binds with
 * [DllImport("KutuphaneCL", CallingConvention = Cdecl)], plus a small set of `ck*`
 * extension entry points used by the C++ host layer (libcekirdekler_host.so).
 *
 * Every declaration cites the reference P/Invoke site it replaces; paths are relative to
 * /root/reference/Cekirdekler/Cekirdekler/.  The reference's native side (KutuphaneCL.dll,
 * C++ over OpenCL) is not in the reference tree, so the contract is the C# side.
 *
 * Conventions:
 *  - every object is an opaque handle returned by createX and released by deleteX;
 *    deletes may arrive late, in any order and from any thread: children keep their parents
 *    alive by reference count and deleting an unknown/already-deleted handle is a no-op;
 *  - C# `bool` is marshalled as a 4-byte BOOL: declared ck_bool (int32_t) here;
 *  - errors never throw: int-returning calls give 0 for success, and the text of the
 *    last error on the calling thread is available through ckGetLastError();
 *  - all sizes/offsets/ranges are 32-bit ELEMENT counts; byte math is 64-bit inside;
 *  - every export is thread-safe and selects its CUDA device from the handle it is given.
 */
#ifndef KUTUPHANECL_H
#define KUTUPHANECL_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void*   ck_handle;
typedef int32_t ck_bool;

/* ---- platform list -------------------------------------------------- Cores.cs:43-49 */
ck_handle platformList(void);
int32_t   numberOfPlatforms(ck_handle hList);
void      deletePlatformList(ck_handle hList);

/* ---- platform ------------------------------------------------------- ClPlatform.cs:35-95 */
ck_handle createPlatform(ck_handle hList, int32_t index);
void      deletePlatform(ck_handle hPlatform);
ck_handle getPlatformInfo(ck_handle hPlatform);              /* imported, never called */
ck_handle getPlatformNameString(ck_handle hPlatform);        /* string owned by the platform */
ck_handle getPlatformVendorNameString(ck_handle hPlatform);  /* string owned by the platform */
int32_t   CODE_CPU(void);
int32_t   CODE_GPU(void);
int32_t   CODE_ACC(void);
int32_t   numberOfCpusInPlatform(ck_handle hPlatform);
int32_t   numberOfGpusInPlatform(ck_handle hPlatform);
int32_t   numberOfAcceleratorsInPlatform(ck_handle hPlatform);

/* ---- device --------------------------------------------------------- ClDevice.cs:31-53 */
ck_handle createDevice(ck_handle hPlatform, int32_t type_, int32_t index_);
ck_handle createDeviceAsPartition(ck_handle hPlatform, int32_t type_, int32_t index_, int32_t numberOfCpuCores_);
void      deleteDevice(ck_handle hDevice);
void      getDeviceName(ck_handle hDevice, ck_handle hString);
void      getDeviceVendorName(ck_handle hDevice, ck_handle hString);
ck_bool   deviceGDDR(ck_handle hDevice);
int32_t   deviceComputeUnits(ck_handle hDevice);
uint64_t  deviceMemSize(ck_handle hDevice);

/* ---- context -------------------------------------------------------- ClContext.cs:32-37 */
ck_handle createContext(ck_handle hPlatform, ck_handle hDevice);
void      deleteContext(ck_handle hContext);

/* ---- command queue (= one CUDA stream) ------------------------------ ClCommandQueue.cs:31-47 */
ck_handle createCommandQueue(ck_handle hContext, ck_handle hDevice, int32_t async_);
ck_handle createCommandQueue2(ck_handle hContext, ck_handle hDevice, int32_t async_, ck_bool defaultQueue);
void      deleteCommandQueue(ck_handle hCommandQueue);
void      addMarkerToCommandQueue(ck_handle hCommandQueue);
int32_t   getMarkerCounterOfCommandQueue(ck_handle hCommandQueue);
void      resetMarkerCounterOfCommandQueue(ck_handle hCommandQueue);

/* ---- event / event array / user event ------------------------------- ClEvent.cs:30-34, ClEventArray.cs:30-38, ClUserEvent.cs:30-46 */
ck_handle createEvent(void);
void      deleteEvent(ck_handle hEvent);
ck_handle createEventArr(ck_bool cpy);
void      addToEventArr(ck_handle hArr, ck_handle hEvent, ck_bool isCopy);
void      deleteEventArr(ck_handle hArr);
ck_handle createUserEvent(ck_handle hContext);
void      deleteUserEvent(ck_handle hUserEvent);
void      triggerUserEvent(ck_handle hUserEvent);
void      addUserEvent(ck_handle hCommandQueue, ck_handle hUserEvent);
void      decrementUserEvent(ck_handle hUserEvent, ck_handle hContext);
void      incrementUserEvent(ck_handle hUserEvent);

/* ---- string --------------------------------------------------------- ClString.cs:30-37, JsonCPPCS.cs:32-40 */
ck_handle   createString(void);
ck_handle   writeToString(ck_handle hString, const char* text);
void        deleteString(ck_handle hString);
const char* readFromString(ck_handle hStringInfo);      /* borrowed pointer */
void        jsonStringCallBack(ck_handle hStringInfo);  /* "read done": releases the string */

/* ---- program (NVRTC → sm_100a cubin) -------------------------------- ClProgram.cs:30-40 */
ck_handle createProgram(ck_handle hContext, ck_handle hDevice, ck_handle hString);
void      deleteProgram(ck_handle hProgram);
int32_t   getProgramErr(ck_handle hProgram);             /* 0 = built */
char*     readProgramErrString(ck_handle hProgram);      /* malloc'd copy; caller frees */

/* ---- kernel --------------------------------------------------------- ClKernel.cs:31-38, Worker.cs:50, Cores.cs:40 */
ck_handle createKernel(ck_handle hProgram, ck_handle hString);
void      deleteKernel(ck_handle hKernel);
int32_t   getKernelErr(ck_handle hKernel);               /* 0 = found */
void      setKernelArgument(ck_handle hKernel, ck_handle hBuffer, int32_t index);
void      setKernelArguments(ck_handle hKernel, ck_handle hBuffer, int32_t index_); /* imported, never called */

/* ---- nd-range (a boxed int: offset, global or local size) ----------- ClNdRange.cs:31-35 */
ck_handle createNdRange(int32_t n);
void      deleteNdRange(ck_handle hRange);

/* ---- launch --------------------------------------------------------- Worker.cs:36-46 */
int32_t compute(ck_handle hCommandQueue, ck_handle hKernel, ck_handle hGlobalRangeReference, ck_handle hGlobalRange, ck_handle hLocalRange);
int32_t computeRepeated(ck_handle hCommandQueue, ck_handle hKernel, ck_handle hGlobalRangeReference, ck_handle hGlobalRange, ck_handle hLocalRange, int32_t repeats);
int32_t computeRepeatedWithSyncKernel(ck_handle hCommandQueue, ck_handle hKernel, ck_handle hGlobalRangeReference, ck_handle hGlobalRange, ck_handle hLocalRange, int32_t repeats, ck_handle hSyncKernel, ck_handle hZeroRange);
int32_t computeEvent(ck_handle hCommandQueue, ck_handle hKernel, ck_handle hGlobalRangeReference, ck_handle hGlobalRange, ck_handle hLocalRange, ck_handle hEventArr, ck_handle hEvt);

/* ---- synchronisation ------------------------------------------------ Worker.cs:52-65, Cores.cs:39-49 */
void finish(ck_handle hCommandQueue);
void flush(ck_handle hCommandQueue);
void wait2(ck_handle hCommandQueue, ck_handle hCommandQueue2);
void wait3(ck_handle hCommandQueue, ck_handle hCommandQueue2, ck_handle hCommandQueue3);
void waitN(ck_handle* hCommandQueueListToWait, ck_handle hCommandQueueToSync, int32_t nQueues);

/* ---- device buffer -------------------------------------------------- ClBuffer.cs:32-35, :259-260 */
/* clArrayType: ClBuffer.SizeOf — float 0, double 1, int 2, long 3, half 4 (C# char, 2 B), char 5 (byte), uint 6 */
ck_handle createBuffer(ck_handle hContext, int32_t numberOfElements, int32_t clArrayType, int32_t isCSharpArray,
                       void* arrayPointer, ck_bool GDDR_BUFFER, ck_bool readOnly, ck_bool writeOnly);
void      deleteBuffer(ck_handle hBuffer);

/* ---- transfers: every typed C# overload collapses onto one void* symbol ---- ClBuffer.cs:37-256 */
/* "write" = host→device, "read" = device→host; ranged forms move ELEMENTS [reference, reference+range)
 * at the same element offset on both sides. */
ck_handle writeToBuffer(ck_handle hCommandQueue, ck_handle hBuffer, void* hostArray);
ck_handle writeToBufferEvent(ck_handle hCommandQueue, ck_handle hBuffer, void* hostArray, ck_handle hEventArr, ck_handle hEvt);
ck_handle writeToBufferRanged(ck_handle hCommandQueue, ck_handle hBuffer, int32_t reference, int32_t range, void* hostArray);
ck_handle writeToBufferRangedEvent(ck_handle hCommandQueue, ck_handle hBuffer, int32_t reference, int32_t range, void* hostArray, ck_handle hEventArr, ck_handle hEvt);
ck_handle readFromBuffer(ck_handle hCommandQueue, ck_handle hBuffer, void* hostArray);
ck_handle readFromBufferEvent(ck_handle hCommandQueue, ck_handle hBuffer, void* hostArray, ck_handle hEventArr, ck_handle hEvt);
ck_handle readFromBufferRanged(ck_handle hCommandQueue, ck_handle hBuffer, int32_t reference, int32_t range, void* hostArray);
ck_handle readFromBufferRangedEvent(ck_handle hCommandQueue, ck_handle hBuffer, int32_t reference, int32_t range, void* hostArray, ck_handle hEventArr, ck_handle hEvt);

/* ---- host ("C space") arrays: pinned, mapped, aligned --------------- CSpaceArrays.cs:108-147 */
/* typeOfArray: CSpaceArrays.ARR_* — FLOAT 0, DOUBLE 1, LONG 2, INT 3, UINT 4, BYTE 5, CHAR 6 (2 B) */
void      sizeOf(int32_t* definitionArr /* int[20] */);
ck_handle createArray(int32_t numberOfElements, int32_t alignment, int32_t typeOfArray);
void*     alignedArrHead(ck_handle hArr);
void      deleteArray(ck_handle hArr);
void      copyMemory(void* dest, void* src, uint32_t count);

/* ======================================================================================
 * Extensions (no counterpart in the reference).  Used by libcekirdekler_host.so and tests.
 * ====================================================================================== */

/* Text of the last error raised on the calling thread ("" if none); borrowed pointer. */
const char* ckGetLastError(void);
/* Number of kernels this library has launched since load (process-wide). */
uint64_t    ckLaunchCount(void);
/* Number of CUDA devices visible, or 0 when no driver/GPU is present (never fails). */
int32_t     ckDeviceCount(void);

/* Device-side timing for the load balancer (replaces the reference's host Stopwatch,
 * Worker.cs:760-807).  A timer is a pair of timing-enabled events. */
ck_handle ckCreateTimer(ck_handle hContext);
void      ckDeleteTimer(ck_handle hTimer);
void      ckTimerStart(ck_handle hTimer, ck_handle hCommandQueue);
void      ckTimerStop(ck_handle hTimer, ck_handle hCommandQueue);
double    ckTimerElapsedMs(ck_handle hTimer);   /* blocks until the stop event completes; <0 on error */

/* Raw views for peer copies and collectives. */
void*     ckBufferDevicePointer(ck_handle hBuffer);
int64_t   ckBufferBytes(ck_handle hBuffer);
int32_t   ckBufferIsZeroCopy(ck_handle hBuffer);
void*     ckQueueStream(ck_handle hCommandQueue);           /* cudaStream_t */
int32_t   ckContextDeviceOrdinal(ck_handle hContext);

/* Device-to-device hand-off over NVLink P2P: copies ELEMENTS [reference, reference+range) of src into the
 * same element range of dst (buffers may live on different devices), ordered on hCommandQueue after the
 * events in hEventArr (may be NULL) and signalling hEvt (may be NULL).  Returns 0 on success. */
int32_t   ckCopyBufferPeer(ck_handle hCommandQueue, ck_handle hDstBuffer, ck_handle hSrcBuffer,
                           int32_t reference, int32_t range, ck_handle hEventArr, ck_handle hEvt);
/* Whether `deviceA` can map `deviceB`'s memory (1) or copies will be staged (0). */
int32_t   ckPeerAccess(int32_t deviceA, int32_t deviceB);

/* Make a queue wait for an event / record an event on a queue (building blocks for hand-off). */
void      ckQueueWaitEvent(ck_handle hCommandQueue, ck_handle hEvent);
void      ckQueueRecordEvent(ck_handle hCommandQueue, ck_handle hEvent);
/* Host-side wait for an event (no-op if never recorded). */
void      ckEventSynchronize(ck_handle hEvent);

/* Page-lock caller memory (e.g. a C# T[] or numpy array) so transfers stay asynchronous. */
int32_t   ckHostRegister(void* ptr, uint64_t bytes);
int32_t   ckHostUnregister(void* ptr);

/* Fill a device buffer's bytes with zero on a queue (used for histogram outputs). */
int32_t   ckZeroBuffer(ck_handle hCommandQueue, ck_handle hBuffer);

/* Whole-array broadcast of a `read` input that every device needs: uploads the host array once
 * over the host link into one buffer (the root rotates between calls), then ncclBroadcast carries it to the other
 * buffers over NVLink, each device's part ordered on its own queue.  When NCCL cannot be loaded (or CK_BROADCAST=p2p)
 * a binary tree of peer copies from buffers[0] does the same job.  All buffers must have the same byte size and live on
 * distinct devices; queues[i] is a queue on buffers[i]'s device.  Returns 0 on success. */
int32_t   ckBroadcastUpload(ck_handle* hQueues, ck_handle* hBuffers, int32_t nDevices, void* hostArray);
/* How many broadcasts went through NCCL and through peer-copy trees since the library was loaded. */
void      ckBroadcastCounts(uint64_t* viaNccl, uint64_t* viaPeerCopies);

/* The same broadcast when every GPU has its own PROCESS (torchrun): a NCCL communicator over the ranks of the job.
 * Rank 0 calls ckSpmdUniqueId (128 bytes out) and passes the id to the others through the job's own channel; every rank
 * then calls ckSpmdInit (collective).  ckSpmdBroadcast sends the whole buffer of rank `root` into the same buffer of every
 * other rank, in place, ordered on each rank's queue; every rank must call it, with the same root.  All return 0 on
 * success and a negative value when NCCL cannot be loaded (the caller then uploads per rank as the reference does). */
int32_t   ckSpmdUniqueId(void* out128);
int32_t   ckSpmdInit(const void* id128, int32_t rank, int32_t world, int32_t deviceOrdinal);
int32_t   ckSpmdWorld(void);      /* ranks in the communicator, 0 before ckSpmdInit */
int32_t   ckSpmdBroadcast(ck_handle hCommandQueue, ck_handle hBuffer, int32_t root);
void      ckSpmdFinalize(void);

/* Device buffers of other processes (one process per GPU), for kernels that store into every GPU's copy of an array over
 * NVLink: export a 64-byte handle of a cudaMalloc'd buffer of this process, open another process's handle (the returned
 * device pointer is usable by kernels on `device`), close it again.  The owner keeps the buffer alive until every importer
 * has closed it.  Export/close return 0 on success; open returns NULL on failure (ckGetLastError has the text). */
int32_t   ckIpcExport(int32_t device, void* devicePointer, void* out64);
void*     ckIpcOpen(int32_t device, const void* handle64);
int32_t   ckIpcClose(int32_t device, void* importedPointer);

/* One n-body step (the nBody3Dw kernel: a += m d / (|d|^2 + eps2)^1.5 over all n bodies, v += a dt, posOut = p + v dt)
 * for bodies [firstBody, firstBody + bodies), WITHOUT allocating anything.  This one has no counterpart in the reference:
 * there every kernel argument is a buffer made by createBuffer, and making a buffer is a cudaMalloc, which drains every
 * queue of the GPU and can fail.  It is for callers that cannot accept either — code that runs while an enqueue-mode
 * region, a pipeline or a device pool keeps the same GPU's queues busy, and hosts whose device memory already belongs to
 * another allocator.  The kernel's work queue and partial sums live in a 16 MiB array inside the library's device image
 * (held by every GPU that loads the library, until exit) instead of a caller-made scratch buffer.
 *   pos, vel, posOut   DEVICE pointers to n float4 each (e.g. ckBufferDevicePointer); pos is read whole, vel and posOut
 *                      are read/written for the share only.  stream is a cudaStream_t (e.g. ckQueueStream) on `device`.
 *   limits             bodies: a positive multiple of 32, at most 1,048,576 per call; n at most 16,777,216.
 *   returns            0 enqueued (the call does not wait), 1 bad argument, 2 beyond the limits (nothing launched),
 *                      3 CUDA error; ckGetLastError() has the text.
 *   first call         on each device looks up the array's address; make it before starting any stream capture.
 *   sharing            every call on a device uses the same array and clears the counters it needs first, in stream
 *                      order.  Calls whose kernels may overlap on the GPU (different streams, no event between them)
 *                      race on it: keep a device's calls on one stream or order them with events. */
int32_t   ckNBody3DStepNoAlloc(int32_t device, void* stream, const void* pos, void* vel, void* posOut, int32_t n,
                               int32_t firstBody, int32_t bodies, float dt, float eps2);

#ifdef __cplusplus
}
#endif
#endif /* KUTUPHANECL_H */
