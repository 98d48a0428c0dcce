// mandelbrot.cu — K1 of the hot path: escape-time Mandelbrot, FP32-FMA-bound.
//
// The reference tree has no Mandelbrot source (its mandelbrot_bench ships only as a binary: SURVEY.md F4), so the
// arithmetic is defined here and, identically, in oracle/oracle.cpp::oracle_mandelbrot:
//   view = { x0, y0, dx, dy, (float)W, (float)maxIter, -, - }      (scalars travel in a small `read` array: F9)
//   pixel p: px = p % W, py = p / W, cx = fma(px, dx, x0), cy = fma(py, dy, y0), z0 = 0
//   out[p] = first k in [0, maxIter) with fma(x_k, x_k, y_k*y_k) > 4, else maxIter
//   x_{k+1} = fma(x_k, x_k, fma(-y_k, y_k, cx)),  y_{k+1} = fma(2, x_k*y_k, cy)
// Every operation is an explicit round-to-nearest intrinsic, so the compiler can neither fuse nor reorder them and the
// iteration counts are bit-identical to the CPU oracle.
//
// Launch contract (ClNumberCruncher): one work-item = MANDEL_E consecutive pixels (`numberOfElementsPerWorkItem = 4`
// on the output array), so a 256-thread work-group owns 1024 consecutive pixels and a device's share is a contiguous
// pixel range starting at get_global_id-derived offsets.  Each lane stores its 4 counts with one 16-byte store.
//
// Roofline: 4 FMA-pipe operations per iteration (FFMA, FMUL, FFMA-with-immediate, FFMA), none of which reads more than
// two distinct registers.  That matters on Blackwell: an instruction that needs two source registers from the same
// (even/odd) register bank issues over two cycles, and the textbook y' = fma(x+x, y, cy) — three distinct registers —
// always does.  Measured on B200 with all lanes in the set (profiles/experiments micro-benchmark, round 1): textbook 5-op form
// 6.35 issue slots per iteration (63 % of the FP32 peak in 8-FLOP units), this form 4.29 slots (93 %) with blocks of 32.
// The escape test is hoisted out of the inner loop: a block of iterations runs unchecked, then one test; if the orbit left
// the radius-2 disc inside the block, the block is replayed from its saved start with a per-iteration test to find the
// exact count.  That is exact because for |c|^2 <= 3.9 an orbit that has left the disc keeps growing
// (|z'| >= |z|^2 - |c| >= 1.05 |z| once |z| > 2 - 3e-7, far above float rounding error) until it overflows to inf/NaN,
// and !(m <= 4) catches all three; pixels with |c|^2 > 3.9 (a thin ring that escapes within a few iterations, plus the
// needle tip near c = -2) take the per-iteration path from the start.  Work is counted as 8 FLOP per executed iteration
// (SURVEY.md §8d), so the issue ceiling of this formulation is 8 / (2 * (4 + overhead)) of the FFMA peak.

#define MANDEL_E 4
#define MANDEL_K1 8    // short unchecked blocks first: most pixels outside the set leave within a few iterations, and a
#define MANDEL_B1 4    // short block keeps both the wasted unchecked work and the replay small for them ...
#define MANDEL_K2 64   // ... then long blocks, where the per-block test/bookkeeping is 12 instructions per 256 useful ones
// Measured on the full 16384^2 x 4096 image on B200 (profiles/experiments/mv.cu, round 1): one block length of 32 -> 25.31 ms (78.8 % of
// the FP32 peak), 64 -> 25.03 ms, 4 blocks of 8 then blocks of 64 -> 24.04 ms (83.0 %); identical iteration counts.

// exact per-iteration loop: at most `limit` iterations from state (x, y); returns the number executed before escape
__device__ __forceinline__ int mandel_exact(float& x, float& y, const float cx, const float cy, const int limit)
{
    int k = 0;
#pragma unroll 1
    while (k < limit) {
        const float m = __fmaf_rn(x, x, __fmul_rn(y, y));
        if (m > 4.0f) break;
        const float t = __fmaf_rn(-y, y, cx);
        const float w = __fmul_rn(x, y);
        y = __fmaf_rn(2.0f, w, cy);
        x = __fmaf_rn(x, x, t);
        k++;
    }
    return k;
}

// One unchecked block of K iterations followed by one escape test.  Returns true when the orbit left the disc somewhere in
// (n, n + K]; (x, y) is then restored to the block's start so the caller can replay it with the exact test (the test at the
// block's last state is included, so the replay may run all K iterations: escape exactly at n + K).
template <int K>
__device__ __forceinline__ bool mandel_block(float& x, float& y, const float cx, const float cy)
{
    const float xs = x, ys = y;
#pragma unroll
    for (int u = 0; u < K; u++) {
        const float t = __fmaf_rn(-y, y, cx);
        const float w = __fmul_rn(x, y);
        y = __fmaf_rn(2.0f, w, cy);
        x = __fmaf_rn(x, x, t);
    }
    const float m = __fmaf_rn(x, x, __fmul_rn(y, y));
    if (!(m <= 4.0f)) {
        x = xs;
        y = ys;
        return true;
    }
    return false;
}

__device__ __forceinline__ int mandel_pixel(const float cx, const float cy, const int maxIter)
{
    float x = 0.0f, y = 0.0f;
    int n = 0;
    const float c2 = __fmaf_rn(cx, cx, __fmul_rn(cy, cy));
    if (c2 <= 3.9f) {
#pragma unroll 1
        for (int b = 0; b < MANDEL_B1 && n + MANDEL_K1 <= maxIter; b++) {
            if (mandel_block<MANDEL_K1>(x, y, cx, cy)) return n + mandel_exact(x, y, cx, cy, MANDEL_K1);
            n += MANDEL_K1;
        }
#pragma unroll 1
        for (int blocks = (maxIter - n) / MANDEL_K2; blocks > 0; blocks--) {
            if (mandel_block<MANDEL_K2>(x, y, cx, cy)) return n + mandel_exact(x, y, cx, cy, MANDEL_K2);
            n += MANDEL_K2;
        }
    }
    return n + mandel_exact(x, y, cx, cy, maxIter - n);
}

__kernel void mandelbrot(const float* __restrict__ view, int* __restrict__ out)
{
    const float x0 = __ldg(view + 0), y0 = __ldg(view + 1), dx = __ldg(view + 2), dy = __ldg(view + 3);
    const unsigned int W = (unsigned int)__ldg(view + 4);
    const int maxIter = (int)__ldg(view + 5);

    const unsigned int item = (unsigned int)get_global_id(0);
    const unsigned int p0 = item * MANDEL_E;  // first pixel of this work-item; W*H <= 2^31 (C# int ranges, F10)
    int r[MANDEL_E];
#pragma unroll
    for (int e = 0; e < MANDEL_E; e++) {
        const unsigned int p = p0 + e;
        const unsigned int py = p / W, px = p - py * W;
        const float cx = __fmaf_rn((float)px, dx, x0);
        const float cy = __fmaf_rn((float)py, dy, y0);
        r[e] = mandel_pixel(cx, cy, maxIter);
    }
    *reinterpret_cast<int4*>(out + p0) = make_int4(r[0], r[1], r[2], r[3]);
}

// One pixel per work-item (numberOfElementsPerWorkItem = 1): the form the benchmark uses from round 2 on.
// A work-item of `mandelbrot` runs its four pixels one after the other, so a work-group inside the set lasts 4 x maxIter
// iterations (0.5 ms at 4096); when a launch runs out of work-groups every SM idles until its slowest one ends — 2 % of a
// 24 ms launch, but 7 % of the 3.3 ms share one of eight GPUs gets — and a warp covers 128 pixels of one row, the longest
// cut through the set's boundary (29.9 of 32 lanes active per instruction).  Here the serial length of a work-group is
// maxIter iterations and a warp covers 32 pixels.  Measured on B200, 16384^2 x 4096 (profiles/r2_mandelbrot_kernels_1gpu.md):
// mandelbrot 24.05 ms (82.9 % of the FP32 peak), this kernel 23.02 ms (86.7 %); a variant that also gave every warp an
// 8 x 4 pixel footprint inside 32 x 8 tiles was slower (23.61 ms) and was dropped.
__kernel void mandelbrot1(const float* __restrict__ view, int* __restrict__ out)
{
    const float x0 = __ldg(view + 0), y0 = __ldg(view + 1), dx = __ldg(view + 2), dy = __ldg(view + 3);
    const unsigned int W = (unsigned int)__ldg(view + 4);
    const int maxIter = (int)__ldg(view + 5);
    const unsigned int p = (unsigned int)get_global_id(0);
    const unsigned int py = p / W, px = p - py * W;
    out[p] = mandel_pixel(__fmaf_rn((float)px, dx, x0), __fmaf_rn((float)py, dy, y0), maxIter);
}
