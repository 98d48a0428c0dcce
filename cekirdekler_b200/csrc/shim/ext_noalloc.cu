// ext_noalloc.cu — ckNBody3DStepNoAlloc: one n-body step (the nBody3Dw kernel) that allocates nothing.
//
// Why it exists.  Everything else in this library gets device memory by creating a buffer, and creating a buffer calls
// cudaMalloc.  cudaMalloc waits for the whole device: every queue of every cruncher on that GPU drains before it returns.
// Some callers of this code cannot afford that, or cannot allocate at all:
//   * code running while an enqueue-mode region, a ClPipeline or a device pool has transfers and kernels in flight on the
//     same GPU — a new scratch array there stalls the sixteen queues it was written to keep busy;
//   * hosts whose device memory is already handed out in full to another allocator (a framework's caching pool, a MIG
//     slice with a hard quota), where cudaMalloc simply fails;
//   * callers that must not fail half-way through a step for lack of memory.
// nBody3Dw is the one kernel here that needs memory beyond its inputs and outputs (its unit queue and partial sums), so it
// gets a variant whose scratch is a fixed array inside this library's device image.
//
// What to know before calling it.
//   * The scratch is a file-scope `static __device__` array of CKNA_SCRATCH_BYTES (16 MiB + 256 B).  Every GPU on which
//     this library's device code gets loaded holds that much until the process exits, whether or not the entry point is used.
//   * The first call on each device looks the array's device address up (cudaGetSymbolAddress, with that device current)
//     and keeps it; that first call should be made before any stream capture begins.  Later calls only launch two kernels.
//   * All calls on one device share the array, and it keeps what the last call left.  Each call clears the counters it
//     needs to start from zero (in stream order, before its own kernel).  Two calls whose kernels can overlap on the GPU —
//     different streams without an event between them — race on the array and produce garbage: keep the calls of one device
//     on one stream, or order them with events.
//   * A share larger than CKNA_MAX_BODIES bodies, or n above CKNA_MAX_N, is refused (return 2) — nothing is launched.
//   * 16 MiB hold fewer partial sums than nBody3Dw's preferred chunking wants once a call covers more than about 21,000
//     bodies x 131,072 (the kernel then widens its chunks; results stay within the same 1e-4 tolerance but are no longer
//     bit-identical to the array-backed kernel).
// No cudaMalloc, cudaFree, or library routine that allocates is called from here.
#include "handles.h"

#include <cstdint>
#include <cstdio>
#include <mutex>
#include <string>

// ---- the kernel source is the shipped one, compiled statically: supply what the run-time rewrite would have added ----
#define __kernel static __global__
#define get_global_id(d) ((size_t)(__ck_goff + blockIdx.x * blockDim.x + threadIdx.x))
#define get_global_offset(d) ((size_t)(__ck_goff))
#define nBody2D(...) ckna_nBody2D(const unsigned int __ck_goff, __VA_ARGS__)
#define nBody3D1(...) ckna_nBody3D1(const unsigned int __ck_goff, __VA_ARGS__)
#define nBody3D(...) ckna_nBody3D(const unsigned int __ck_goff, __VA_ARGS__)
#define nBody3Dq(...) ckna_nBody3Dq(const unsigned int __ck_goff, __VA_ARGS__)
#define nBody3Dw(...) ckna_nBody3Dw(const unsigned int __ck_goff, __VA_ARGS__)
#define nBody3Dwp(...) ckna_nBody3Dwp(const unsigned int __ck_goff, __VA_ARGS__)
#pragma nv_diag_suppress 177   // the file's other kernels come along unused
#include "../kernels/nbody.cu"
#undef nBody3Dwp
#undef nBody2D
#undef nBody3D1
#undef nBody3D
#undef nBody3Dq
#undef nBody3Dw
#undef get_global_id
#undef get_global_offset
#undef __kernel

#define CKNA_PARAM_FLOATS 64                        // head of the array: the kernel's params block (256 bytes)
#define CKNA_SCRATCH_KFLOATS 4096                   // the kernel's own scratch: 4096 x 1024 floats = 16 MiB
#define CKNA_SCRATCH_FLOATS (CKNA_SCRATCH_KFLOATS * 1024)
#define CKNA_SCRATCH_BYTES ((CKNA_PARAM_FLOATS + CKNA_SCRATCH_FLOATS) * 4)
#define CKNA_MAX_BODIES 1048576                     // per call: 8192 groups x 384 floats = 12 MiB of partials at one chunk
#define CKNA_MAX_N 16777216                         // n travels to the kernel as a float
#define CKNA_WORK_ITEMS_PER_BODY 4
#define CKNA_LOCAL 128
#define CKNA_MAX_DEVICES 64

// In host code this name is only a handle for cudaGetSymbolAddress — it is not a device address.
static __device__ __align__(256) float ckna_scratch[CKNA_PARAM_FLOATS + CKNA_SCRATCH_FLOATS];

// Runs before the step's kernel, in stream order: writes the params block and clears the queue header and the per-group
// counters (the part of the shared array that must start at zero; the partial sums are overwritten before they are read).
static __global__ void ckna_prepare(float* base, float dt, float eps2, int n, int counters)
{
    int* ctr = reinterpret_cast<int*>(base + CKNA_PARAM_FLOATS);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < counters; i += gridDim.x * blockDim.x) ctr[i] = 0;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        base[0] = dt;
        base[1] = eps2;
        base[2] = (float)n;
        base[3] = (float)CKNA_WORK_ITEMS_PER_BODY;
        base[4] = (float)CKNA_SCRATCH_KFLOATS;
    }
}

namespace {
std::mutex g_mu;
float* g_base[CKNA_MAX_DEVICES] = {};   // device address of ckna_scratch, per device ordinal

// once per device, with that device current
float* scratch_on(int device)
{
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_base[device]) {
        float* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), ckna_scratch);
        if (e != cudaSuccess || !p) {
            ck::set_error(std::string("ckNBody3DStepNoAlloc: cudaGetSymbolAddress on device ") + std::to_string(device) + ": " +
                          cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")");
            return nullptr;
        }
        g_base[device] = p;
    }
    return g_base[device];
}
}  // namespace

extern "C" int32_t ckNBody3DStepNoAlloc(int32_t device, void* stream, const void* pos, void* vel, void* posOut, int32_t n,
                                        int32_t firstBody, int32_t bodies, float dt, float eps2)
{
    if (device < 0 || device >= CKNA_MAX_DEVICES || !pos || !vel || !posOut || n <= 0 || firstBody < 0 || bodies <= 0 ||
        (long long)firstBody + bodies > n || bodies % (CKNA_LOCAL / CKNA_WORK_ITEMS_PER_BODY) != 0) {
        ck::set_error("ckNBody3DStepNoAlloc: bad argument (device pointers must be non-null, 0 <= firstBody, firstBody + bodies <= n, "
                      "bodies a positive multiple of 32)");
        return 1;
    }
    if (bodies > CKNA_MAX_BODIES || n > CKNA_MAX_N) {
        ck::set_error("ckNBody3DStepNoAlloc: " + std::to_string(bodies) + " bodies of " + std::to_string(n) +
                      " exceed what the built-in scratch serves (" + std::to_string(CKNA_MAX_BODIES) + " bodies per call, n <= " +
                      std::to_string(CKNA_MAX_N) + "); use nBody3Dw with a scratch array of your own");
        return 2;
    }
    int previous = -1;
    cudaGetDevice(&previous);
    if (!ck::cuda_ok(cudaSetDevice(device), "ckNBody3DStepNoAlloc: cudaSetDevice")) return 3;
    int32_t rc = 0;
    float* base = scratch_on(device);
    if (!base) {
        rc = 3;
    } else {
        cudaStream_t s = static_cast<cudaStream_t>(stream);
        const int groups = (bodies + NBW_GROUP - 1) / NBW_GROUP;
        const int counters = NBW_HDR + groups;
        // The runtime keeps the last error of ANY earlier call on this thread until somebody reads it (a library probing for
        // optional driver entry points leaves cudaErrorSymbolNotFound behind, seen on an 8-GPU box after NCCL start-up):
        // drop that first, so that what is read after each launch below belongs to the launch.
        (void)cudaGetLastError();
        ckna_prepare<<<(counters + 255) / 256, 256, 0, s>>>(base, dt, eps2, n, counters);
        if (!ck::cuda_ok(cudaGetLastError(), "ckNBody3DStepNoAlloc: launch of the prepare kernel")) rc = 3;
        if (rc == 0) {
            const unsigned int grid = (unsigned int)(bodies / (CKNA_LOCAL / CKNA_WORK_ITEMS_PER_BODY));
            ckna_nBody3Dw<<<grid, CKNA_LOCAL, 0, s>>>((unsigned int)firstBody * CKNA_WORK_ITEMS_PER_BODY,
                                                       static_cast<const float4*>(pos), static_cast<float4*>(vel),
                                                       static_cast<float4*>(posOut), base, base + CKNA_PARAM_FLOATS);
            if (!ck::cuda_ok(cudaGetLastError(), "ckNBody3DStepNoAlloc: launch of nBody3Dw")) rc = 3;
        }
    }
    if (previous >= 0 && previous != device) cudaSetDevice(previous);
    return rc;
}
