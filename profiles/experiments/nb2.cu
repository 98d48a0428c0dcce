// n-body: B bodies per thread, B j-slices per work-group (1 thread per body overall), packed f32x2 over pairs of j.
#include <cstdio>
#include <cuda_runtime.h>
#include <vector>
#include <cmath>
typedef unsigned long long u64;
#define TILE 256
__device__ __forceinline__ float rsq(float v){ float r; asm("rsqrt.approx.ftz.f32 %0, %1;":"=f"(r):"f"(v)); return r; }
__device__ __forceinline__ u64 pk(float a, float b){ u64 r; asm("mov.b64 %0, {%1,%2};":"=l"(r):"f"(a),"f"(b)); return r;}
__device__ __forceinline__ void upk(u64 v, float&a, float&b){ asm("mov.b64 {%0,%1}, %2;":"=f"(a),"=f"(b):"l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a,u64 b,u64 c){ u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;":"=l"(d):"l"(a),"l"(b),"l"(c)); return d;}
__device__ __forceinline__ u64 mul2(u64 a,u64 b){ u64 d; asm("mul.rn.f32x2 %0, %1, %2;":"=l"(d):"l"(a),"l"(b)); return d;}
__device__ __forceinline__ u64 sub2(u64 a,u64 b){ u64 d; asm("sub.rn.f32x2 %0, %1, %2;":"=l"(d):"l"(a),"l"(b)); return d;}

template <int B, int UNROLL>
__global__ void __launch_bounds__(256) nbk(const float4* __restrict__ pos, float4* __restrict__ acc, int n, float eps2, int firstBody) {
    // block of L threads owns L bodies; B slices of L/B threads; thread q of a slice carries bodies q*B .. q*B+B-1 of the block
    __shared__ float4 sA[B][TILE / 2];
    __shared__ float4 sB[B][TILE / 2];
    __shared__ float red[3 * B][256];
    const int LS = blockDim.x / B;
    const int slice = threadIdx.x / LS, q = threadIdx.x - slice * LS;
    const int blockFirst = firstBody + blockIdx.x * blockDim.x;
    u64 PX_[B], PY_[B], PZ_[B], AX[B], AY[B], AZ[B];
#pragma unroll
    for (int b = 0; b < B; b++) {
        const float4 p = pos[blockFirst + q * B + b];
        PX_[b] = pk(p.x, p.x); PY_[b] = pk(p.y, p.y); PZ_[b] = pk(p.z, p.z);
        AX[b] = pk(0.f, 0.f); AY[b] = AX[b]; AZ[b] = AX[b];
    }
    const u64 EPS = pk(eps2, eps2);
    const int tiles = (n + TILE - 1) / TILE;
    const int tilesPerSlice = (tiles + B - 1) / B;
    const int firstTile = slice * tilesPerSlice;
    float* fA = reinterpret_cast<float*>(sA[slice]);
    float* fB = reinterpret_cast<float*>(sB[slice]);
    for (int it = 0; it < tilesPerSlice; it++) {
        const int tile = firstTile + it;
        for (int t = q; t < TILE; t += LS) {
            const int j = tile * TILE + t;
            const float4 pj = (tile < tiles && j < n) ? pos[j] : make_float4(0.f, 0.f, 0.f, 0.f);
            const int p = t >> 1, h = t & 1;
            fA[p * 4 + h] = pj.x; fA[p * 4 + 2 + h] = pj.y; fB[p * 4 + h] = pj.z; fB[p * 4 + 2 + h] = pj.w;
        }
        __syncthreads();
        const float4* tA = sA[slice];
        const float4* tB = sB[slice];
#pragma unroll UNROLL
        for (int t = 0; t < TILE / 2; t++) {
            const float4 a = tA[t], bq = tB[t];
            const u64 PX = pk(a.x, a.y), PY = pk(a.z, a.w), PZ = pk(bq.x, bq.y), PM = pk(bq.z, bq.w);
#pragma unroll
            for (int b = 0; b < B; b++) {
                const u64 DX = sub2(PX, PX_[b]), DY = sub2(PY, PY_[b]), DZ = sub2(PZ, PZ_[b]);
                const u64 R2 = fma2(DZ, DZ, fma2(DY, DY, fma2(DX, DX, EPS)));
                float r0, r1; upk(R2, r0, r1);
                const u64 INV = pk(rsq(r0), rsq(r1));
                const u64 S = mul2(PM, mul2(mul2(INV, INV), INV));
                AX[b] = fma2(DX, S, AX[b]); AY[b] = fma2(DY, S, AY[b]); AZ[b] = fma2(DZ, S, AZ[b]);
            }
        }
        __syncthreads();
    }
    // combine the B slices through shared memory: slice s > 0 writes, slice 0 sums
    float ax[B], ay[B], az[B];
#pragma unroll
    for (int b = 0; b < B; b++) { float lo, hi; upk(AX[b], lo, hi); ax[b] = lo + hi; upk(AY[b], lo, hi); ay[b] = lo + hi; upk(AZ[b], lo, hi); az[b] = lo + hi; }
    for (int s = 1; s < B; s++) {
        if (slice == s) {
#pragma unroll
            for (int b = 0; b < B; b++) { red[3 * b][q] = ax[b]; red[3 * b + 1][q] = ay[b]; red[3 * b + 2][q] = az[b]; }
        }
        __syncthreads();
        if (slice == 0) {
#pragma unroll
            for (int b = 0; b < B; b++) { ax[b] += red[3 * b][q]; ay[b] += red[3 * b + 1][q]; az[b] += red[3 * b + 2][q]; }
        }
        __syncthreads();
    }
    if (slice == 0) {
#pragma unroll
        for (int b = 0; b < B; b++) acc[blockFirst + q * B + b] = make_float4(ax[b], ay[b], az[b], 0.f);
    }
}

int main() {
    int n = 131072; std::vector<float4> h(n); unsigned s = 12345; auto rnd = [&]() { s = s * 1664525u + 1013904223u; return (s >> 8) * (1.0f / 16777216.0f) * 30.f - 15.f; };
    for (auto& p : h) { p.x = rnd(); p.y = rnd(); p.z = rnd(); p.w = 1.f; }
    float4 *d, *acc; cudaMalloc(&d, n * 16); cudaMalloc(&acc, n * 16); cudaMemcpy(d, h.data(), n * 16, cudaMemcpyHostToDevice);
    int sms, clk; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0); cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    std::vector<float4> ref(n), out(n);
    struct V { const char* name; int B, L, U; } vs[] = {{"B2 L128 u2", 2, 128, 2}, {"B2 L256 u2", 2, 256, 2}, {"B4 L128 u1", 4, 128, 1}, {"B4 L256 u1", 4, 256, 1}, {"B4 L128 u2", 4, 128, 2}, {"B4 L256 u2", 4, 256, 2}};
    for (int share : {131072, 16384}) {
        for (auto& v : vs) {
            const int grid = share / v.L;
            auto launch = [&]() {
                if (v.B == 2 && v.U == 2) nbk<2, 2><<<grid, v.L>>>(d, acc, n, 1e-4f, 0);
                else if (v.B == 4 && v.U == 1) nbk<4, 1><<<grid, v.L>>>(d, acc, n, 1e-4f, 0);
                else nbk<4, 2><<<grid, v.L>>>(d, acc, n, 1e-4f, 0);
            };
            launch(); cudaDeviceSynchronize();
            cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
            float best = 1e9;
            for (int r = 0; r < 3; r++) { cudaEventRecord(a); launch(); cudaEventRecord(b); cudaEventSynchronize(b); float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms; }
            cudaMemcpy(out.data(), acc, (size_t)share * 16, cudaMemcpyDeviceToHost);
            if (&v == &vs[0] && share == 131072) ref = out;
            double maxrel = 0; for (int i = 0; i < share; i++) { double dd = fabs(out[i].x - ref[i].x) / fmax(1e-2, fabs(ref[i].x)); if (dd > maxrel) maxrel = dd; }
            double inter = (double)share * n; double ginter = inter / best / 1e6; double frac = ginter * 1e9 * 20 / (2.0 * 128 * sms * clk * 1e3);
            printf("share %6d %-12s %7.3f ms %8.1f GInter/s  frac %.3f  maxrel %.1e  %s\n", share, v.name, best, ginter, frac, maxrel, cudaGetErrorString(cudaGetLastError()));
        }
    }
    return 0;
}
