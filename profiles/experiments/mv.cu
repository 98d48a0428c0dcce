// Mandelbrot kernel variants on the real C3 workload (16384^2 x 4096): time + checksum (sum of counts must match).
#include <cstdio>
#include <cuda_runtime.h>
#include <vector>

__device__ __forceinline__ int exact(float& x, float& y, float cx, float cy, int limit) {
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
#define STEP { const float t = __fmaf_rn(-y, y, cx); const float w = __fmul_rn(x, y); y = __fmaf_rn(2.0f, w, cy); x = __fmaf_rn(x, x, t); }

// V0: production (K=32, while n+K<=maxIter)
template <int K> __device__ __forceinline__ int pix0(float cx, float cy, int maxIter) {
    float x = 0.f, y = 0.f; int n = 0;
    const float c2 = __fmaf_rn(cx, cx, __fmul_rn(cy, cy));
    if (c2 <= 3.9f) {
        while (n + K <= maxIter) {
            const float xs = x, ys = y;
#pragma unroll
            for (int u = 0; u < K; u++) STEP
            const float m = __fmaf_rn(x, x, __fmul_rn(y, y));
            if (!(m <= 4.0f)) { x = xs; y = ys; return n + exact(x, y, cx, cy, K); }
            n += K;
        }
    }
    return n + exact(x, y, cx, cy, maxIter - n);
}
// V1: exact prologue of P iterations, then counted blocks of K
template <int K, int P> __device__ __forceinline__ int pix1(float cx, float cy, int maxIter) {
    float x = 0.f, y = 0.f;
    const float c2 = __fmaf_rn(cx, cx, __fmul_rn(cy, cy));
    int n = exact(x, y, cx, cy, min(P, maxIter));
    if (n < min(P, maxIter)) return n;                              // escaped inside the checked prologue
    if (c2 > 3.9f) return n + exact(x, y, cx, cy, maxIter - n);     // ring near |c| = 2: per-iteration path
    int blocks = (maxIter - n) / K;
#pragma unroll 1
    for (; blocks > 0; blocks--) {
        const float xs = x, ys = y;
#pragma unroll
        for (int u = 0; u < K; u++) STEP
        const float m = __fmaf_rn(x, x, __fmul_rn(y, y));
        if (!(m <= 4.0f)) { x = xs; y = ys; return n + exact(x, y, cx, cy, K); }
        n += K;
    }
    return n + exact(x, y, cx, cy, maxIter - n);
}
// V2: small first blocks (K1 x B1) then big blocks K2
template <int K1, int B1, int K2> __device__ __forceinline__ int pix2(float cx, float cy, int maxIter) {
    float x = 0.f, y = 0.f; int n = 0;
    const float c2 = __fmaf_rn(cx, cx, __fmul_rn(cy, cy));
    if (c2 <= 3.9f) {
#pragma unroll 1
        for (int b = 0; b < B1 && n + K1 <= maxIter; b++) {
            const float xs = x, ys = y;
#pragma unroll
            for (int u = 0; u < K1; u++) STEP
            const float m = __fmaf_rn(x, x, __fmul_rn(y, y));
            if (!(m <= 4.0f)) { x = xs; y = ys; return n + exact(x, y, cx, cy, K1); }
            n += K1;
        }
        int blocks = (maxIter - n) / K2;
#pragma unroll 1
        for (; blocks > 0; blocks--) {
            const float xs = x, ys = y;
#pragma unroll
            for (int u = 0; u < K2; u++) STEP
            const float m = __fmaf_rn(x, x, __fmul_rn(y, y));
            if (!(m <= 4.0f)) { x = xs; y = ys; return n + exact(x, y, cx, cy, K2); }
            n += K2;
        }
    }
    return n + exact(x, y, cx, cy, maxIter - n);
}

#define KERNEL(name, CALL)                                                                              \
    extern "C" __global__ void name(const float* view, int* out) {                                      \
        const float x0 = view[0], y0 = view[1], dx = view[2], dy = view[3];                             \
        const unsigned W = (unsigned)view[4]; const int maxIter = (int)view[5];                         \
        const unsigned p0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4; int r[4];                      \
        _Pragma("unroll") for (int e = 0; e < 4; e++) {                                                 \
            const unsigned p = p0 + e, py = p / W, px = p - py * W;                                     \
            const float cx = __fmaf_rn((float)px, dx, x0), cy = __fmaf_rn((float)py, dy, y0);           \
            r[e] = CALL;                                                                                \
        }                                                                                               \
        *reinterpret_cast<int4*>(out + p0) = make_int4(r[0], r[1], r[2], r[3]);                         \
    }
KERNEL(k_prod32, (pix0<32>(cx, cy, maxIter)))
KERNEL(k_prod64, (pix0<64>(cx, cy, maxIter)))
KERNEL(k_pro8_32, (pix1<32, 8>(cx, cy, maxIter)))
KERNEL(k_pro8_64, (pix1<64, 8>(cx, cy, maxIter)))
KERNEL(k_pro16_64, (pix1<64, 16>(cx, cy, maxIter)))
KERNEL(k_8x2_64, (pix2<8, 2, 64>(cx, cy, maxIter)))
KERNEL(k_8x4_64, (pix2<8, 4, 64>(cx, cy, maxIter)))
KERNEL(k_16x2_64, (pix2<16, 2, 64>(cx, cy, maxIter)))
KERNEL(k_8x4_128, (pix2<8, 4, 128>(cx, cy, maxIter)))

typedef void (*kern)(const float*, int*);
__global__ void sumk(const int* a, size_t n, unsigned long long* s) {
    unsigned long long t = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) t += a[i];
    atomicAdd(s, t);
}
int main() {
    const int W = 16384, H = 16384, IT = 4096;
    float hv[8] = {-2.f, -1.5f, 3.f / W, 3.f / H, (float)W, (float)IT, 0, 0};
    float* view; int* out; unsigned long long* s;
    cudaMalloc(&view, 32); cudaMalloc(&out, (size_t)W * H * 4); cudaMalloc(&s, 8);
    cudaMemcpy(view, hv, 32, cudaMemcpyHostToDevice);
    struct { const char* n; kern k; } ks[] = {{"prod K32", k_prod32}, {"prod K64", k_prod64}, {"exact8+K32", k_pro8_32}, {"exact8+K64", k_pro8_64},
                                              {"exact16+K64", k_pro16_64}, {"8x2 then 64", k_8x2_64}, {"8x4 then 64", k_8x4_64}, {"16x2 then 64", k_16x2_64}, {"8x4 then 128", k_8x4_128}};
    int sms, clk; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0); cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    for (auto& e : ks) {
        const int grid = W * H / 4 / 256;
        e.k<<<grid, 256>>>(view, out); cudaDeviceSynchronize();
        cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
        float best = 1e9;
        for (int r = 0; r < 3; r++) { cudaEventRecord(a); e.k<<<grid, 256>>>(view, out); cudaEventRecord(b); cudaEventSynchronize(b); float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms; }
        cudaMemset(s, 0, 8); sumk<<<1024, 256>>>(out, (size_t)W * H, s); unsigned long long hs; cudaMemcpy(&hs, s, 8, cudaMemcpyDeviceToHost);
        double frac = 8.0 * hs / (best * 1e-3) / (2.0 * 128 * sms * clk * 1e3);
        printf("%-14s %8.3f ms  sum %llu  Mpix/s %9.1f  frac %.4f  %s\n", e.n, best, hs, (double)W * H / best / 1e3, frac, cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
