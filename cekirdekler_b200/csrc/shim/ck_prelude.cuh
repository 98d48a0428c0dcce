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
#define get_group_id(d)      ((size_t)(blockIdx.x))   /* as in OpenCL: 0 .. get_num_groups-1 of THIS enqueue (a device's share) */
#define ck_global_group_id() ((size_t)(__ck_goff / blockDim.x + blockIdx.x))   /* extension: group index within the whole range */
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

// ---- vector types: component-wise operators, literals and geometric built-ins (OpenCL C 6.4, 6.12.5) ----
// OpenCL's `(float4)(a, b, c, d)` literal is rewritten by the host side to make_float4(a, b, c, d); the overloads below add
// the forms CUDA's make_* lack: one scalar replicated, an existing vector, and vectors mixed with scalars.
#define CK_VEC2_OPS(V, S, OP)                                                                                            \
    __device__ __forceinline__ V operator OP(V a, V b) { return make_##V(a.x OP b.x, a.y OP b.y); }                      \
    __device__ __forceinline__ V operator OP(V a, S s) { return make_##V(a.x OP s, a.y OP s); }                          \
    __device__ __forceinline__ V operator OP(S s, V a) { return make_##V(s OP a.x, s OP a.y); }                          \
    __device__ __forceinline__ V& operator OP##=(V& a, V b) { a.x OP##= b.x; a.y OP##= b.y; return a; }                  \
    __device__ __forceinline__ V& operator OP##=(V& a, S s) { a.x OP##= s; a.y OP##= s; return a; }
#define CK_VEC3_OPS(V, S, OP)                                                                                            \
    __device__ __forceinline__ V operator OP(V a, V b) { return make_##V(a.x OP b.x, a.y OP b.y, a.z OP b.z); }          \
    __device__ __forceinline__ V operator OP(V a, S s) { return make_##V(a.x OP s, a.y OP s, a.z OP s); }                \
    __device__ __forceinline__ V operator OP(S s, V a) { return make_##V(s OP a.x, s OP a.y, s OP a.z); }                \
    __device__ __forceinline__ V& operator OP##=(V& a, V b) { a.x OP##= b.x; a.y OP##= b.y; a.z OP##= b.z; return a; }   \
    __device__ __forceinline__ V& operator OP##=(V& a, S s) { a.x OP##= s; a.y OP##= s; a.z OP##= s; return a; }
#define CK_VEC4_OPS(V, S, OP)                                                                                                        \
    __device__ __forceinline__ V operator OP(V a, V b) { return make_##V(a.x OP b.x, a.y OP b.y, a.z OP b.z, a.w OP b.w); }          \
    __device__ __forceinline__ V operator OP(V a, S s) { return make_##V(a.x OP s, a.y OP s, a.z OP s, a.w OP s); }                  \
    __device__ __forceinline__ V operator OP(S s, V a) { return make_##V(s OP a.x, s OP a.y, s OP a.z, s OP a.w); }                  \
    __device__ __forceinline__ V& operator OP##=(V& a, V b) { a.x OP##= b.x; a.y OP##= b.y; a.z OP##= b.z; a.w OP##= b.w; return a; } \
    __device__ __forceinline__ V& operator OP##=(V& a, S s) { a.x OP##= s; a.y OP##= s; a.z OP##= s; a.w OP##= s; return a; }
#define CK_VEC_FAMILY(B)                                                                                                 \
    CK_VEC2_OPS(B##2, B, +) CK_VEC2_OPS(B##2, B, -) CK_VEC2_OPS(B##2, B, *) CK_VEC2_OPS(B##2, B, /)                      \
    CK_VEC3_OPS(B##3, B, +) CK_VEC3_OPS(B##3, B, -) CK_VEC3_OPS(B##3, B, *) CK_VEC3_OPS(B##3, B, /)                      \
    CK_VEC4_OPS(B##4, B, +) CK_VEC4_OPS(B##4, B, -) CK_VEC4_OPS(B##4, B, *) CK_VEC4_OPS(B##4, B, /)                      \
    __device__ __forceinline__ B##2 operator-(B##2 a) { return make_##B##2(-a.x, -a.y); }                                \
    __device__ __forceinline__ B##3 operator-(B##3 a) { return make_##B##3(-a.x, -a.y, -a.z); }                          \
    __device__ __forceinline__ B##4 operator-(B##4 a) { return make_##B##4(-a.x, -a.y, -a.z, -a.w); }                    \
    __device__ __forceinline__ B##2 make_##B##2(B s) { return make_##B##2(s, s); }                                       \
    __device__ __forceinline__ B##3 make_##B##3(B s) { return make_##B##3(s, s, s); }                                    \
    __device__ __forceinline__ B##4 make_##B##4(B s) { return make_##B##4(s, s, s, s); }                                 \
    __device__ __forceinline__ B##2 make_##B##2(B##2 v) { return v; }                                                    \
    __device__ __forceinline__ B##3 make_##B##3(B##3 v) { return v; }                                                    \
    __device__ __forceinline__ B##4 make_##B##4(B##4 v) { return v; }                                                    \
    __device__ __forceinline__ B##3 make_##B##3(B##2 v, B z) { return make_##B##3(v.x, v.y, z); }                        \
    __device__ __forceinline__ B##3 make_##B##3(B x, B##2 v) { return make_##B##3(x, v.x, v.y); }                        \
    __device__ __forceinline__ B##4 make_##B##4(B##2 v, B z, B w) { return make_##B##4(v.x, v.y, z, w); }                \
    __device__ __forceinline__ B##4 make_##B##4(B x, B##2 v, B w) { return make_##B##4(x, v.x, v.y, w); }                \
    __device__ __forceinline__ B##4 make_##B##4(B x, B y, B##2 v) { return make_##B##4(x, y, v.x, v.y); }                \
    __device__ __forceinline__ B##4 make_##B##4(B##2 a, B##2 b) { return make_##B##4(a.x, a.y, b.x, b.y); }              \
    __device__ __forceinline__ B##4 make_##B##4(B##3 v, B w) { return make_##B##4(v.x, v.y, v.z, w); }                   \
    __device__ __forceinline__ B##4 make_##B##4(B x, B##3 v) { return make_##B##4(x, v.x, v.y, v.z); }
CK_VEC_FAMILY(float)
CK_VEC_FAMILY(double)
CK_VEC_FAMILY(int)
CK_VEC_FAMILY(uint)

#define CK_GEOMETRY(B, SQRT, RSQRT)                                                                                      \
    __device__ __forceinline__ B dot(B##2 a, B##2 b) { return a.x * b.x + a.y * b.y; }                                   \
    __device__ __forceinline__ B dot(B##3 a, B##3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }                       \
    __device__ __forceinline__ B dot(B##4 a, B##4 b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }           \
    __device__ __forceinline__ B length(B##2 a) { return SQRT(dot(a, a)); }                                              \
    __device__ __forceinline__ B length(B##3 a) { return SQRT(dot(a, a)); }                                              \
    __device__ __forceinline__ B length(B##4 a) { return SQRT(dot(a, a)); }                                              \
    __device__ __forceinline__ B distance(B##2 a, B##2 b) { return length(a - b); }                                      \
    __device__ __forceinline__ B distance(B##3 a, B##3 b) { return length(a - b); }                                      \
    __device__ __forceinline__ B distance(B##4 a, B##4 b) { return length(a - b); }                                      \
    __device__ __forceinline__ B##2 normalize(B##2 a) { return a * RSQRT(dot(a, a)); }                                   \
    __device__ __forceinline__ B##3 normalize(B##3 a) { return a * RSQRT(dot(a, a)); }                                   \
    __device__ __forceinline__ B##4 normalize(B##4 a) { return a * RSQRT(dot(a, a)); }                                   \
    __device__ __forceinline__ B##3 cross(B##3 a, B##3 b) {                                                              \
        return make_##B##3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);                         \
    }                                                                                                                    \
    __device__ __forceinline__ B##4 cross(B##4 a, B##4 b) {                                                              \
        return make_##B##4(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x, (B)0);                   \
    }
CK_GEOMETRY(float, sqrtf, rsqrtf)
CK_GEOMETRY(double, sqrt, rsqrt)

#define CK_MAP1(F, G)                                                                                                    \
    __device__ __forceinline__ float2 F(float2 a) { return make_float2(G(a.x), G(a.y)); }                                \
    __device__ __forceinline__ float3 F(float3 a) { return make_float3(G(a.x), G(a.y), G(a.z)); }                        \
    __device__ __forceinline__ float4 F(float4 a) { return make_float4(G(a.x), G(a.y), G(a.z), G(a.w)); }
#define CK_MAP2(F, G)                                                                                                    \
    __device__ __forceinline__ float2 F(float2 a, float2 b) { return make_float2(G(a.x, b.x), G(a.y, b.y)); }            \
    __device__ __forceinline__ float3 F(float3 a, float3 b) { return make_float3(G(a.x, b.x), G(a.y, b.y), G(a.z, b.z)); } \
    __device__ __forceinline__ float4 F(float4 a, float4 b) { return make_float4(G(a.x, b.x), G(a.y, b.y), G(a.z, b.z), G(a.w, b.w)); }
CK_MAP1(sqrt, sqrtf) CK_MAP1(rsqrt, rsqrtf) CK_MAP1(fabs, fabsf) CK_MAP1(floor, floorf) CK_MAP1(ceil, ceilf)
CK_MAP1(exp, expf) CK_MAP1(log, logf) CK_MAP1(sin, sinf) CK_MAP1(cos, cosf) CK_MAP1(native_sqrt, __fsqrt_rn)
CK_MAP1(native_rsqrt, rsqrtf) CK_MAP1(native_exp, __expf) CK_MAP1(native_log, __logf) CK_MAP1(native_sin, __sinf)
CK_MAP1(native_cos, __cosf)
CK_MAP2(fmin, fminf) CK_MAP2(fmax, fmaxf) CK_MAP2(min, fminf) CK_MAP2(max, fmaxf) CK_MAP2(pow, powf)
__device__ __forceinline__ float2 mad(float2 a, float2 b, float2 c) { return make_float2(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y)); }
__device__ __forceinline__ float3 mad(float3 a, float3 b, float3 c) { return make_float3(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y), fmaf(a.z, b.z, c.z)); }
__device__ __forceinline__ float4 mad(float4 a, float4 b, float4 c) { return make_float4(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y), fmaf(a.z, b.z, c.z), fmaf(a.w, b.w, c.w)); }
__device__ __forceinline__ float4 fma(float4 a, float4 b, float4 c) { return mad(a, b, c); }
__device__ __forceinline__ float2 clamp(float2 v, float lo, float hi) { return make_float2(clamp(v.x, lo, hi), clamp(v.y, lo, hi)); }
__device__ __forceinline__ float3 clamp(float3 v, float lo, float hi) { return make_float3(clamp(v.x, lo, hi), clamp(v.y, lo, hi), clamp(v.z, lo, hi)); }
__device__ __forceinline__ float4 clamp(float4 v, float lo, float hi) { return make_float4(clamp(v.x, lo, hi), clamp(v.y, lo, hi), clamp(v.z, lo, hi), clamp(v.w, lo, hi)); }
__device__ __forceinline__ float2 mix(float2 a, float2 b, float t) { return a + (b - a) * t; }
__device__ __forceinline__ float3 mix(float3 a, float3 b, float t) { return a + (b - a) * t; }
__device__ __forceinline__ float4 mix(float4 a, float4 b, float t) { return a + (b - a) * t; }
__device__ __forceinline__ float4 convert_float4(int4 v) { return make_float4((float)v.x, (float)v.y, (float)v.z, (float)v.w); }
__device__ __forceinline__ float4 convert_float4(uint4 v) { return make_float4((float)v.x, (float)v.y, (float)v.z, (float)v.w); }
__device__ __forceinline__ int4   convert_int4(float4 v) { return make_int4((int)v.x, (int)v.y, (int)v.z, (int)v.w); }
__device__ __forceinline__ float2 convert_float2(int2 v) { return make_float2((float)v.x, (float)v.y); }
__device__ __forceinline__ int2   convert_int2(float2 v) { return make_int2((int)v.x, (int)v.y); }

// ---- vloadN / vstoreN (OpenCL C 6.12.7): element offset is in units of the vector ----
__device__ __forceinline__ float2 vload2(size_t i, const float* p) { return make_float2(p[2 * i], p[2 * i + 1]); }
__device__ __forceinline__ float3 vload3(size_t i, const float* p) { return make_float3(p[3 * i], p[3 * i + 1], p[3 * i + 2]); }
__device__ __forceinline__ float4 vload4(size_t i, const float* p) { return make_float4(p[4 * i], p[4 * i + 1], p[4 * i + 2], p[4 * i + 3]); }
__device__ __forceinline__ void vstore2(float2 v, size_t i, float* p) { p[2 * i] = v.x; p[2 * i + 1] = v.y; }
__device__ __forceinline__ void vstore3(float3 v, size_t i, float* p) { p[3 * i] = v.x; p[3 * i + 1] = v.y; p[3 * i + 2] = v.z; }
__device__ __forceinline__ void vstore4(float4 v, size_t i, float* p) { p[4 * i] = v.x; p[4 * i + 1] = v.y; p[4 * i + 2] = v.z; p[4 * i + 3] = v.w; }

// ---- integer built-ins (OpenCL C 6.12.3) ----
__device__ __forceinline__ int  mul24(int a, int b) { return __mul24(a, b); }
__device__ __forceinline__ int  mad24(int a, int b, int c) { return __mul24(a, b) + c; }
__device__ __forceinline__ uint mul24(uint a, uint b) { return __umul24(a, b); }
__device__ __forceinline__ uint mad24(uint a, uint b, uint c) { return __umul24(a, b) + c; }
__device__ __forceinline__ int  mul_hi(int a, int b) { return __mulhi(a, b); }
__device__ __forceinline__ uint mul_hi(uint a, uint b) { return __umulhi(a, b); }
__device__ __forceinline__ int  popcount(uint x) { return __popc(x); }
__device__ __forceinline__ int  clz(uint x) { return __clz((int)x); }
__device__ __forceinline__ uint rotate(uint x, uint r) { return __funnelshift_l(x, x, r); }
__device__ __forceinline__ uint abs_diff(int a, int b) { return a > b ? (uint)a - (uint)b : (uint)b - (uint)a; }
__device__ __forceinline__ float step(float edge, float x) { return x < edge ? 0.0f : 1.0f; }
__device__ __forceinline__ float smoothstep(float e0, float e1, float x) {
    const float t = clamp((x - e0) / (e1 - e0), 0.0f, 1.0f);
    return t * t * (3.0f - 2.0f * t);
}
__device__ __forceinline__ float fract(float x) { return fminf(x - floorf(x), 0x1.fffffep-1f); }
__device__ __forceinline__ float degrees(float r) { return r * 57.29577951308232f; }
__device__ __forceinline__ float radians(float d) { return d * 0.017453292519943295f; }
template <typename T> __device__ __forceinline__ T select(T a, T b, int c) { return c ? b : a; }

// ---- component selections beyond .x .y .z .w (OpenCL C 6.1.7) ----
// The host side rewrites `e.xy`, `e.zw`, `e.lo`, `e.hi`, `e.xyz` into ck_xy(e) ... (and `.s0`..`.s3` into `.x`..`.w`).  On an
// lvalue the result is a pair/triple of references that converts to the narrower vector and can be assigned one; on an
// rvalue it is the narrower vector itself.  A struct of the user's own that has a member called lo, hi, xy ... keeps it.
template <class V2, class S> struct ck_sel2 {
    S& a; S& b;
    __device__ __forceinline__ operator V2() const { V2 r; r.x = a; r.y = b; return r; }
    __device__ __forceinline__ ck_sel2& operator=(const V2& v) { a = v.x; b = v.y; return *this; }
    __device__ __forceinline__ ck_sel2& operator=(const ck_sel2& o) { const S t0 = o.a, t1 = o.b; a = t0; b = t1; return *this; }
};
template <class V3, class S> struct ck_sel3 {
    S& a; S& b; S& c;
    __device__ __forceinline__ operator V3() const { V3 r; r.x = a; r.y = b; r.z = c; return r; }
    __device__ __forceinline__ ck_sel3& operator=(const V3& v) { a = v.x; b = v.y; c = v.z; return *this; }
};
#define CK_SWIZZLES(B)                                                                                                        \
    __device__ __forceinline__ ck_sel2<B##2, B> ck_xy(B##4& v) { return {v.x, v.y}; }                                         \
    __device__ __forceinline__ ck_sel2<B##2, B> ck_zw(B##4& v) { return {v.z, v.w}; }                                         \
    __device__ __forceinline__ ck_sel2<B##2, B> ck_lo(B##4& v) { return {v.x, v.y}; }                                         \
    __device__ __forceinline__ ck_sel2<B##2, B> ck_hi(B##4& v) { return {v.z, v.w}; }                                         \
    __device__ __forceinline__ ck_sel3<B##3, B> ck_xyz(B##4& v) { return {v.x, v.y, v.z}; }                                   \
    __device__ __forceinline__ ck_sel2<B##2, B> ck_xy(B##3& v) { return {v.x, v.y}; }                                         \
    __device__ __forceinline__ B##2 ck_xy(const B##4& v) { return make_##B##2(v.x, v.y); }                                    \
    __device__ __forceinline__ B##2 ck_zw(const B##4& v) { return make_##B##2(v.z, v.w); }                                    \
    __device__ __forceinline__ B##2 ck_lo(const B##4& v) { return make_##B##2(v.x, v.y); }                                    \
    __device__ __forceinline__ B##2 ck_hi(const B##4& v) { return make_##B##2(v.z, v.w); }                                    \
    __device__ __forceinline__ B##3 ck_xyz(const B##4& v) { return make_##B##3(v.x, v.y, v.z); }                              \
    __device__ __forceinline__ B##2 ck_xy(const B##3& v) { return make_##B##2(v.x, v.y); }                                    \
    __device__ __forceinline__ B& ck_lo(B##2& v) { return v.x; }                                                              \
    __device__ __forceinline__ B& ck_hi(B##2& v) { return v.y; }                                                              \
    __device__ __forceinline__ B ck_lo(const B##2& v) { return v.x; }                                                         \
    __device__ __forceinline__ B ck_hi(const B##2& v) { return v.y; }
CK_SWIZZLES(float)
CK_SWIZZLES(double)
CK_SWIZZLES(int)
CK_SWIZZLES(uint)
// fall-backs: user structs with members of these names
template <class T> __device__ __forceinline__ auto ck_xy(T&& v) -> decltype((v.xy)) { return v.xy; }
template <class T> __device__ __forceinline__ auto ck_zw(T&& v) -> decltype((v.zw)) { return v.zw; }
template <class T> __device__ __forceinline__ auto ck_lo(T&& v) -> decltype((v.lo)) { return v.lo; }
template <class T> __device__ __forceinline__ auto ck_hi(T&& v) -> decltype((v.hi)) { return v.hi; }
template <class T> __device__ __forceinline__ auto ck_xyz(T&& v) -> decltype((v.xyz)) { return v.xyz; }

#endif  // CK_PRELUDE_CUH
