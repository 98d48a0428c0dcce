// ck_prelude.cuh — device-side prelude that libKutuphaneCL.so puts in front of every kernel
// source string before NVRTC compiles it for sm_100a.
//
// The reference hands OpenCL C strings to clBuildProgram (ClProgram.cs:64-67, Worker.cs:274-279);
// the kernels it ships are plain 1-D buffer kernels (Tester.cs:35-38, :7727-7742, :7820-7824,
// Kamera.cs:234-245).  This prelude lets such strings compile as CUDA C++:
//   * address-space qualifiers become no-ops / __shared__,
//   * work-item functions map onto blockIdx/threadIdx plus the hidden `__ck_goff` parameter
//     (the device's start index: the reference passes it as the NDRange global offset,
//     Worker.cs:1037),
//   * a few OpenCL built-ins and vector operators are provided.
// The host side rewrites `__kernel void f(args)` into
//   extern "C" __global__ void f(const unsigned int __ck_goff, args)
// so `get_global_id(0)` is only usable inside kernel bodies (not in helper functions).
// Hand-written CUDA kernels see the same prelude and may ignore all of it.
#ifndef CK_PRELUDE_CUH
#define CK_PRELUDE_CUH

typedef unsigned int   uint;
typedef unsigned char  uchar;
typedef unsigned short ushort;
typedef unsigned long  ulong;

#define __global
#define __private
#define __constant const
#define __local __shared__
#define __read_only
#define __write_only
#define restrict __restrict__

#define CLK_LOCAL_MEM_FENCE  1
#define CLK_GLOBAL_MEM_FENCE 2

// ---- work-item functions (1-D only: ClNdRange.cs:32) ----
#define get_global_id(d)     ((size_t)(__ck_goff + blockIdx.x * blockDim.x + threadIdx.x))
#define get_global_offset(d) ((size_t)(__ck_goff))
#define get_local_id(d)      ((size_t)(threadIdx.x))
#define get_local_size(d)    ((size_t)(blockDim.x))
#define get_group_id(d)      ((size_t)(__ck_goff / blockDim.x + blockIdx.x))
#define get_num_groups(d)    ((size_t)(gridDim.x))
#define get_global_size(d)   ((size_t)(gridDim.x * blockDim.x))
#define get_work_dim()       (1u)

#define barrier(flags)            __syncthreads()
#define work_group_barrier(flags) __syncthreads()
#define mem_fence(flags)          __threadfence_block()
#define read_mem_fence(flags)     __threadfence_block()
#define write_mem_fence(flags)    __threadfence_block()

// ---- scalar built-ins ----
__device__ __forceinline__ float mad(float a, float b, float c) { return fmaf(a, b, c); }
__device__ __forceinline__ double mad(double a, double b, double c) { return fma(a, b, c); }
__device__ __forceinline__ float native_sqrt(float x) { return __fsqrt_rn(x); }
__device__ __forceinline__ float native_rsqrt(float x) { return rsqrtf(x); }
__device__ __forceinline__ float native_recip(float x) { return __frcp_rn(x); }
__device__ __forceinline__ float native_divide(float a, float b) { return __fdividef(a, b); }
__device__ __forceinline__ float native_exp(float x) { return __expf(x); }
__device__ __forceinline__ float native_log(float x) { return __logf(x); }
__device__ __forceinline__ float native_sin(float x) { return __sinf(x); }
__device__ __forceinline__ float native_cos(float x) { return __cosf(x); }
__device__ __forceinline__ float native_powr(float a, float b) { return __powf(a, b); }
__device__ __forceinline__ float half_sqrt(float x) { return __fsqrt_rn(x); }
__device__ __forceinline__ float half_rsqrt(float x) { return rsqrtf(x); }
__device__ __forceinline__ float clamp(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
__device__ __forceinline__ int   clamp(int x, int lo, int hi) { return min(max(x, lo), hi); }
__device__ __forceinline__ float mix(float a, float b, float t) { return a + (b - a) * t; }
__device__ __forceinline__ float sign(float x) { return (x > 0.0f) ? 1.0f : ((x < 0.0f) ? -1.0f : 0.0f); }
__device__ __forceinline__ int   convert_int(float x) { return (int)x; }
__device__ __forceinline__ uint  convert_uint(float x) { return (uint)x; }
__device__ __forceinline__ float convert_float(int x) { return (float)x; }
__device__ __forceinline__ int   as_int(float x) { return __float_as_int(x); }
__device__ __forceinline__ uint  as_uint(float x) { return __float_as_uint(x); }
__device__ __forceinline__ float as_float(int x) { return __int_as_float(x); }
__device__ __forceinline__ float as_float(uint x) { return __uint_as_float(x); }

// ---- atomics (OpenCL 1.2 spellings; all operate on global or local ints) ----
template <typename T> __device__ __forceinline__ T atomic_add(volatile T* p, T v) { return atomicAdd((T*)p, v); }
template <typename T> __device__ __forceinline__ T atomic_sub(volatile T* p, T v) { return atomicSub((T*)p, v); }
template <typename T> __device__ __forceinline__ T atomic_inc(volatile T* p) { return atomicAdd((T*)p, (T)1); }
template <typename T> __device__ __forceinline__ T atomic_dec(volatile T* p) { return atomicSub((T*)p, (T)1); }
template <typename T> __device__ __forceinline__ T atomic_min(volatile T* p, T v) { return atomicMin((T*)p, v); }
template <typename T> __device__ __forceinline__ T atomic_max(volatile T* p, T v) { return atomicMax((T*)p, v); }
template <typename T> __device__ __forceinline__ T atomic_xchg(volatile T* p, T v) { return atomicExch((T*)p, v); }
template <typename T> __device__ __forceinline__ T atomic_cmpxchg(volatile T* p, T c, T v) { return atomicCAS((T*)p, c, v); }

// ---- float2/float4 component-wise operators (OpenCL vector arithmetic) ----
__device__ __forceinline__ float2 operator+(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 operator-(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 operator*(float2 a, float2 b) { return make_float2(a.x * b.x, a.y * b.y); }
__device__ __forceinline__ float2 operator*(float2 a, float s) { return make_float2(a.x * s, a.y * s); }
__device__ __forceinline__ float2 operator*(float s, float2 a) { return make_float2(a.x * s, a.y * s); }
__device__ __forceinline__ float4 operator+(float4 a, float4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
__device__ __forceinline__ float4 operator-(float4 a, float4 b) { return make_float4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
__device__ __forceinline__ float4 operator*(float4 a, float4 b) { return make_float4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
__device__ __forceinline__ float4 operator*(float4 a, float s) { return make_float4(a.x * s, a.y * s, a.z * s, a.w * s); }
__device__ __forceinline__ float4 operator*(float s, float4 a) { return make_float4(a.x * s, a.y * s, a.z * s, a.w * s); }
__device__ __forceinline__ float4& operator+=(float4& a, float4 b) { a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; return a; }
__device__ __forceinline__ float4& operator-=(float4& a, float4 b) { a.x -= b.x; a.y -= b.y; a.z -= b.z; a.w -= b.w; return a; }
__device__ __forceinline__ float  dot(float4 a, float4 b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
__device__ __forceinline__ float  dot(float2 a, float2 b) { return a.x * b.x + a.y * b.y; }
__device__ __forceinline__ float  length(float4 a) { return sqrtf(dot(a, a)); }
__device__ __forceinline__ float  length(float2 a) { return sqrtf(dot(a, a)); }

#endif  // CK_PRELUDE_CUH
