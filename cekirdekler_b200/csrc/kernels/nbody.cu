// nbody.cu — K4 of the hot path: all-pairs N-body, FP32 + SFU bound, shared-memory tiled.
//
//   nBody2D   the reference's kernel (Tester.cs:7727-7742) with its semantics kept exactly: 2-D, separate x/y
//             arrays, every j including j == i, softening 1e-4 inside the square root, and fy accumulating dx
//             (quirk Q3, present in the reference's host loop too).  The reference splices n into the source; here n
//             travels in a 1-element `read` array so one compiled program serves every n.  Tiled through shared
//             memory so each position is read from HBM/L2 once per work-group instead of once per work-item.
//             Arithmetic is written with explicit non-fused intrinsics in the reference's source order, so results
//             are bit-identical to oracle_nbody2d (sqrtf and the division are IEEE-rounded on both sides).
//   nBody3D   the BASELINE float4 step (no reference source; defined here and in oracle_nbody3d_step):
//             a_i = sum_j m_j d / (|d|^2 + eps2)^(3/2), v += a dt, posOut = p + v dt;  params = { dt, eps2, (float)n }.
//             One work-item = one body (numberOfElementsPerWorkItem = 4 floats on vel/posOut); pos is a whole-array
//             `read` input (every device needs all of it: the one broadcast of the path).  The inverse square root is
//             the raw SFU approximation (MUFU.RSQ); parity with the oracle is <= 1e-4 relative on positions.
// Work is counted as 20 FLOP per interaction.

#define NB_TILE 256   // bodies staged per shared-memory tile; equals the work-group size used by the benchmarks

__device__ __forceinline__ float ck_rsqrt_approx(float v)
{
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v));
    return r;
}

__kernel void nBody2D(const float* __restrict__ x, const float* __restrict__ y, float* __restrict__ fx,
                      float* __restrict__ fy, const int* __restrict__ nArr)
{
    __shared__ float sx[NB_TILE];
    __shared__ float sy[NB_TILE];
    const int n = nArr[0];
    const int i = (int)get_global_id(0);
    const float xi = x[i], yi = y[i];
    float fx0 = 0.0f, fy0 = 0.0f;
    for (int base = 0; base < n; base += NB_TILE) {
        for (int t = threadIdx.x; t < NB_TILE; t += blockDim.x) {
            const int j = base + t;
            sx[t] = (j < n) ? x[j] : 0.0f;
            sy[t] = (j < n) ? y[j] : 0.0f;
        }
        __syncthreads();
        const int lim = min(NB_TILE, n - base);
#pragma unroll 4
        for (int t = 0; t < lim; t++) {
            const float dx = __fsub_rn(xi, sx[t]);
            const float dy = __fsub_rn(yi, sy[t]);
            const float r = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), 0.0001f));
            const float r3 = __fmul_rn(__fmul_rn(r, r), r);
            const float inv = __fdiv_rn(1.0f, r3);
            const float f = __fmul_rn(dx, inv);
            fx0 = __fadd_rn(fx0, f);
            fy0 = __fadd_rn(fy0, f);  // (sic) the reference accumulates dx into fy as well
        }
        __syncthreads();
    }
    fx[i] = fx0;
    fy[i] = fy0;
}

__kernel void nBody3D(const float4* __restrict__ pos, float4* __restrict__ vel, float4* __restrict__ posOut,
                      const float* __restrict__ params)
{
    __shared__ float4 tile[NB_TILE];
    const float dt = __ldg(params + 0), eps2 = __ldg(params + 1);
    const int n = (int)__ldg(params + 2);
    const int i = (int)get_global_id(0);
    const float4 pi = pos[i];
    float ax = 0.0f, ay = 0.0f, az = 0.0f;
    for (int base = 0; base < n; base += NB_TILE) {
        for (int t = threadIdx.x; t < NB_TILE; t += blockDim.x) {
            const int j = base + t;
            tile[t] = (j < n) ? pos[j] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);  // padding bodies have mass 0
        }
        __syncthreads();
#pragma unroll 8
        for (int t = 0; t < NB_TILE; t++) {
            const float4 pj = tile[t];  // same address for the whole warp: one broadcast LDS.128
            const float dx = __fsub_rn(pj.x, pi.x), dy = __fsub_rn(pj.y, pi.y), dz = __fsub_rn(pj.z, pi.z);
            const float r2 = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmaf_rn(dx, dx, eps2)));
            const float inv = ck_rsqrt_approx(r2);
            const float s = __fmul_rn(pj.w, __fmul_rn(__fmul_rn(inv, inv), inv));
            ax = __fmaf_rn(dx, s, ax);
            ay = __fmaf_rn(dy, s, ay);
            az = __fmaf_rn(dz, s, az);
        }
        __syncthreads();
    }
    float4 v = vel[i];
    v.x = __fmaf_rn(ax, dt, v.x);
    v.y = __fmaf_rn(ay, dt, v.y);
    v.z = __fmaf_rn(az, dt, v.z);
    vel[i] = v;
    posOut[i] = make_float4(__fmaf_rn(v.x, dt, pi.x), __fmaf_rn(v.y, dt, pi.y), __fmaf_rn(v.z, dt, pi.z), pi.w);
}
