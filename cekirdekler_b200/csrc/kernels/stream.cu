// stream.cu — streaming kernels K2/K3 of the hot path: HBM-bound when the arrays are resident,
// host-link-bound when ClNumberCruncher streams them with partial reads/writes.
//
//   vectorAdd   c[i] = a[i] + b[i], 1 element per work-item, the operation of the reference's stream test
//               (Tester.cs:7820-7824; its verbatim kernel text is a test vector under tests/reference_kernels.py).
//   vectorAdd4  same result, 4 elements per work-item (numberOfElementsPerWorkItem = 4): one 16-byte load per input
//               and one 16-byte store per warp lane, fully coalesced (512 B per warp per access).
//   saxpy / saxpy4   y[i] = alpha[0]*x[i] + y[i] with the multiply and the add rounded separately
//               (__fmul_rn/__fadd_rn are never contracted into an FMA), so results are bit-identical to the
//               contraction-off CPU oracle.  alpha travels in a 1-element whole-`read` array: the API has no
//               scalar kernel arguments.
// Algorithmic traffic: 12 B per element (two 4 B reads + one 4 B write).

__kernel void vectorAdd(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ c)
{
    const size_t i = get_global_id(0);
    c[i] = __fadd_rn(a[i], b[i]);
}

__kernel void vectorAdd4(const float4* __restrict__ a, const float4* __restrict__ b, float4* __restrict__ c)
{
    const size_t i = get_global_id(0);
    const float4 va = __ldcs(a + i);   // streaming loads: every byte is touched exactly once
    const float4 vb = __ldcs(b + i);
    float4 r;
    r.x = __fadd_rn(va.x, vb.x);
    r.y = __fadd_rn(va.y, vb.y);
    r.z = __fadd_rn(va.z, vb.z);
    r.w = __fadd_rn(va.w, vb.w);
    __stcs(c + i, r);
}

__kernel void saxpy(const float* __restrict__ alpha, const float* __restrict__ x, float* __restrict__ y)
{
    const size_t i = get_global_id(0);
    y[i] = __fadd_rn(__fmul_rn(alpha[0], x[i]), y[i]);
}

__kernel void saxpy4(const float* __restrict__ alpha, const float4* __restrict__ x, float4* __restrict__ y)
{
    const size_t i = get_global_id(0);
    const float a = __ldg(alpha);
    const float4 vx = __ldcs(x + i);
    float4 vy = __ldcs(y + i);
    vy.x = __fadd_rn(__fmul_rn(a, vx.x), vy.x);
    vy.y = __fadd_rn(__fmul_rn(a, vx.y), vy.y);
    vy.z = __fadd_rn(__fmul_rn(a, vx.z), vy.z);
    vy.w = __fadd_rn(__fmul_rn(a, vx.w), vy.w);
    __stcs(y + i, vy);
}
