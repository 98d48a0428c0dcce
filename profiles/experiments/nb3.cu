// n-body with a device-side work queue: units = (128-body group, j chunk), partials in scratch, last arriver integrates.
#include <cstdio>
#include <cuda_runtime.h>
#include <vector>
#include <cmath>
typedef unsigned long long u64;
#define NB_TILE 256
#define NB_B 4
#define NB_PREFETCH 8
__device__ __forceinline__ float ck_rsqrt_approx(float v){ float r; asm("rsqrt.approx.ftz.f32 %0, %1;":"=f"(r):"f"(v)); return r; }
typedef u64 ck_f32x2;
__device__ __forceinline__ u64 ck_pack(float a, float b){ u64 r; asm("mov.b64 %0, {%1,%2};":"=l"(r):"f"(a),"f"(b)); return r;}
__device__ __forceinline__ void ck_unpack(u64 v, float&a, float&b){ asm("mov.b64 {%0,%1}, %2;":"=f"(a),"=f"(b):"l"(v)); }
__device__ __forceinline__ u64 ck_fma2(u64 a,u64 b,u64 c){ u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;":"=l"(d):"l"(a),"l"(b),"l"(c)); return d;}
__device__ __forceinline__ u64 ck_mul2(u64 a,u64 b){ u64 d; asm("mul.rn.f32x2 %0, %1, %2;":"=l"(d):"l"(a),"l"(b)); return d;}
__device__ __forceinline__ u64 ck_sub2(u64 a,u64 b){ u64 d; asm("sub.rn.f32x2 %0, %1, %2;":"=l"(d):"l"(a),"l"(b)); return d;}
#define NB_PAIR_INTERACT(PIX, PIY, PIZ, AX, AY, AZ)                                          \
    {                                                                                        \
        const ck_f32x2 DX = ck_sub2(PX, PIX), DY = ck_sub2(PY, PIY), DZ = ck_sub2(PZ, PIZ);  \
        const ck_f32x2 R2 = ck_fma2(DZ, DZ, ck_fma2(DY, DY, ck_fma2(DX, DX, EPS)));          \
        float r0, r1;                                                                        \
        ck_unpack(R2, r0, r1);                                                               \
        const ck_f32x2 INV = ck_pack(ck_rsqrt_approx(r0), ck_rsqrt_approx(r1));              \
        const ck_f32x2 S = ck_mul2(PM, ck_mul2(ck_mul2(INV, INV), INV));                     \
        AX = ck_fma2(DX, S, AX);                                                             \
        AY = ck_fma2(DY, S, AY);                                                             \
        AZ = ck_fma2(DZ, S, AZ);                                                             \
    }

#define NB_GROUP 128          // bodies per unit: 32 lanes x NB_B
#define NB_MAXW 8             // warps per work-group at most
#define NB_HDR 4              // ints before the per-group counters: queue, exited, pad, pad

template <int UNROLL, bool PF>
__global__ void nbw(const float4* __restrict__ pos, float4* __restrict__ vel, float4* __restrict__ posOut,
                    const float* __restrict__ params, float* __restrict__ scratch, unsigned goff)
{
    __shared__ float4 sA[NB_MAXW][NB_TILE / 2];
    __shared__ float4 sB[NB_MAXW][NB_TILE / 2];
    __shared__ float red[NB_MAXW - 1][3 * NB_B][32];
    __shared__ int sNext, sLast;
    const float dt = __ldg(params + 0), eps2 = __ldg(params + 1);
    const int n = (int)__ldg(params + 2);
    const long long cap = (long long)__ldg(params + 3);
    const int wpb = (int)__ldg(params + 4);              // work-items per body: 1, 2 or 4
    const int targetUnits = (int)__ldg(params + 5);
    const int S = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int firstBody = (int)(goff / wpb);
    const int share = (int)((long long)gridDim.x * blockDim.x / wpb);
    const int G = (share + NB_GROUP - 1) / NB_GROUP;
    const int tiles = (n + NB_TILE - 1) / NB_TILE;
    int* ctr = reinterpret_cast<int*>(scratch);
    int* done = ctr + NB_HDR;
    const long long P0 = ((long long)(NB_HDR + G) + 127) / 128 * 128;
    // chunks per group: enough units to keep the tail short, at least S tiles per unit, and what scratch can hold
    int C = targetUnits / G; if (C < 1) C = 1;
    { const int cmax = tiles / S; if (C > cmax) C = cmax < 1 ? 1 : cmax; }
    { const long long cfit = (cap - P0) / ((long long)G * 3 * NB_GROUP); if (C > cfit) C = (int)cfit; }
    if (C < 1) { if (blockIdx.x == 0 && threadIdx.x == 0) printf("nBody3D: scratch too small\n"); return; }
    const int chunkTiles = (tiles + C - 1) / C;
    C = (tiles + chunkTiles - 1) / chunkTiles;
    const int totalUnits = G * C;
    float* part = scratch + P0;
    const ck_f32x2 EPS = ck_pack(eps2, eps2);
    float* fA = reinterpret_cast<float*>(sA[warp]);
    float* fB = reinterpret_cast<float*>(sB[warp]);

    if (threadIdx.x == 0) sNext = atomicAdd(&ctr[0], 1);
    __syncthreads();
    int u = sNext;
    while (u < totalUnits) {
        const int g = u / C, c = u - g * C;
        const int lastBody = firstBody + share - 1;
        int i0 = firstBody + g * NB_GROUP + NB_B * lane;
        const bool liveLane = i0 <= lastBody;
        if (!liveLane) i0 = lastBody - (NB_B - 1);
        ck_f32x2 PIX[NB_B], PIY[NB_B], PIZ[NB_B], AX[NB_B], AY[NB_B], AZ[NB_B];
#pragma unroll
        for (int b = 0; b < NB_B; b++) {
            const float4 p = pos[i0 + b];
            PIX[b] = ck_pack(p.x, p.x); PIY[b] = ck_pack(p.y, p.y); PIZ[b] = ck_pack(p.z, p.z);
            AX[b] = ck_pack(0.0f, 0.0f); AY[b] = AX[b]; AZ[b] = AX[b];
        }
        // this warp's tiles of chunk c
        const int chunkFirst = c * chunkTiles;
        int chunkEnd = chunkFirst + chunkTiles; if (chunkEnd > tiles) chunkEnd = tiles;
        const int per = (chunkEnd - chunkFirst + S - 1) / S;
        const int t0 = chunkFirst + warp * per;
        int t1 = t0 + per; if (t1 > chunkEnd) t1 = chunkEnd;
        float4 nxt[NB_PREFETCH];
#pragma unroll
        for (int k = 0; k < NB_PREFETCH; k++) {
            const int j = t0 * NB_TILE + lane + 32 * k;
            if (PF) nxt[k] = (t0 < t1 && j < n) ? pos[j] : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        for (int tile = t0; tile < t1; tile++) {
            if (!PF) {
#pragma unroll
                for (int k = 0; k < NB_PREFETCH; k++) { const int j = tile * NB_TILE + lane + 32 * k; nxt[k] = (j < n) ? pos[j] : make_float4(0.f, 0.f, 0.f, 0.f); }
            }
#pragma unroll
            for (int k = 0; k < NB_PREFETCH; k++) {
                const int t = lane + 32 * k;
                const int p = t >> 1, h = t & 1;
                fA[p * 4 + h] = nxt[k].x; fA[p * 4 + 2 + h] = nxt[k].y; fB[p * 4 + h] = nxt[k].z; fB[p * 4 + 2 + h] = nxt[k].w;
            }
            __syncwarp();
#pragma unroll
            for (int k = 0; k < NB_PREFETCH; k++) {
                const int j = (tile + 1) * NB_TILE + lane + 32 * k;
                if (PF) nxt[k] = (tile + 1 < t1 && j < n) ? pos[j] : make_float4(0.f, 0.f, 0.f, 0.f);
            }
            const float4* tA = sA[warp];
            const float4* tB = sB[warp];
#pragma unroll UNROLL
            for (int t = 0; t < NB_TILE / 2; t++) {
                const float4 a = tA[t], b4 = tB[t];
                const ck_f32x2 PX = ck_pack(a.x, a.y), PY = ck_pack(a.z, a.w), PZ = ck_pack(b4.x, b4.y), PM = ck_pack(b4.z, b4.w);
#pragma unroll
                for (int b = 0; b < NB_B; b++) NB_PAIR_INTERACT(PIX[b], PIY[b], PIZ[b], AX[b], AY[b], AZ[b])
            }
            __syncwarp();
        }
        float acc[3 * NB_B];
#pragma unroll
        for (int b = 0; b < NB_B; b++) {
            float lo, hi;
            ck_unpack(AX[b], lo, hi); acc[3 * b] = lo + hi;
            ck_unpack(AY[b], lo, hi); acc[3 * b + 1] = lo + hi;
            ck_unpack(AZ[b], lo, hi); acc[3 * b + 2] = lo + hi;
        }
        if (warp > 0) {
#pragma unroll
            for (int k = 0; k < 3 * NB_B; k++) red[warp - 1][k][lane] = acc[k];
        }
        __syncthreads();
        if (warp == 0) {
            for (int s = 0; s < S - 1; s++) {
#pragma unroll
                for (int k = 0; k < 3 * NB_B; k++) acc[k] += red[s][k][lane];
            }
            float* dst = part + (long long)u * (3 * NB_GROUP);
#pragma unroll
            for (int b = 0; b < NB_B; b++) {
                dst[0 * NB_GROUP + b * 32 + lane] = acc[3 * b];
                dst[1 * NB_GROUP + b * 32 + lane] = acc[3 * b + 1];
                dst[2 * NB_GROUP + b * 32 + lane] = acc[3 * b + 2];
            }
            __threadfence();
            __syncwarp();
            if (lane == 0) {
                sLast = (atomicAdd(&done[g], 1) == C - 1);
                sNext = atomicAdd(&ctr[0], 1);
            }
        }
        __syncthreads();
        const bool last = sLast != 0;
        const int unext = sNext;
        if (last) {
            __threadfence();
            for (int t = threadIdx.x; t < NB_GROUP; t += blockDim.x) {
                const int body = firstBody + g * NB_GROUP + NB_B * (t & 31) + (t >> 5);
                if (body <= lastBody) {
                    float ax = 0.f, ay = 0.f, az = 0.f;
                    const float* src = part + (long long)g * C * (3 * NB_GROUP) + t;
                    for (int cc = 0; cc < C; cc++) {
                        ax += __ldcg(src + 0 * NB_GROUP); ay += __ldcg(src + 1 * NB_GROUP); az += __ldcg(src + 2 * NB_GROUP);
                        src += 3 * NB_GROUP;
                    }
                    const float4 p = pos[body];
                    float4 v = vel[body];
                    v.x = __fmaf_rn(ax, dt, v.x); v.y = __fmaf_rn(ay, dt, v.y); v.z = __fmaf_rn(az, dt, v.z);
                    vel[body] = v;
                    posOut[body] = make_float4(__fmaf_rn(v.x, dt, p.x), __fmaf_rn(v.y, dt, p.y), __fmaf_rn(v.z, dt, p.z), p.w);
                }
            }
            if (threadIdx.x == 0) done[g] = 0;
        }
        __syncthreads();
        u = unext;
    }
    if (threadIdx.x == 0) {
        if (atomicAdd(&ctr[1], 1) == (int)gridDim.x - 1) { ctr[0] = 0; ctr[1] = 0; }
    }
}

// the static kernel of round 1 for comparison (4 bodies per thread, 4 slices)
__global__ void __launch_bounds__(256) nbk(const float4* __restrict__ pos, float4* __restrict__ vel, float4* __restrict__ posOut, int n, float eps2, float dt, int firstBody) {
    __shared__ float4 sA[NB_B][NB_TILE / 2];
    __shared__ float4 sB[NB_B][NB_TILE / 2];
    __shared__ float red[3 * NB_B][256];
    const int LS = blockDim.x / NB_B;
    const int slice = threadIdx.x / LS, q = threadIdx.x - slice * LS;
    const int blockFirst = firstBody + blockIdx.x * blockDim.x;
    u64 PIX[NB_B], PIY[NB_B], PIZ[NB_B], AX[NB_B], AY[NB_B], AZ[NB_B];
#pragma unroll
    for (int b = 0; b < NB_B; b++) {
        const float4 p = pos[blockFirst + q * NB_B + b];
        PIX[b] = ck_pack(p.x, p.x); PIY[b] = ck_pack(p.y, p.y); PIZ[b] = ck_pack(p.z, p.z);
        AX[b] = ck_pack(0.f, 0.f); AY[b] = AX[b]; AZ[b] = AX[b];
    }
    const u64 EPS = ck_pack(eps2, eps2);
    const int tiles = (n + NB_TILE - 1) / NB_TILE;
    const int tilesPerSlice = (tiles + NB_B - 1) / NB_B;
    const int firstTile = slice * tilesPerSlice;
    float* fA = reinterpret_cast<float*>(sA[slice]);
    float* fB = reinterpret_cast<float*>(sB[slice]);
    for (int it = 0; it < tilesPerSlice; it++) {
        const int tile = firstTile + it;
        for (int t = q; t < NB_TILE; t += LS) {
            const int j = tile * NB_TILE + t;
            const float4 pj = (tile < tiles && j < n) ? pos[j] : make_float4(0.f, 0.f, 0.f, 0.f);
            const int p = t >> 1, h = t & 1;
            fA[p * 4 + h] = pj.x; fA[p * 4 + 2 + h] = pj.y; fB[p * 4 + h] = pj.z; fB[p * 4 + 2 + h] = pj.w;
        }
        __syncthreads();
        const float4* tA = sA[slice];
        const float4* tB = sB[slice];
#pragma unroll 2
        for (int t = 0; t < NB_TILE / 2; t++) {
            const float4 a = tA[t], b4 = tB[t];
            const u64 PX = ck_pack(a.x, a.y), PY = ck_pack(a.z, a.w), PZ = ck_pack(b4.x, b4.y), PM = ck_pack(b4.z, b4.w);
#pragma unroll
            for (int b = 0; b < NB_B; b++) NB_PAIR_INTERACT(PIX[b], PIY[b], PIZ[b], AX[b], AY[b], AZ[b])
        }
        __syncthreads();
    }
    float ax[NB_B], ay[NB_B], az[NB_B];
#pragma unroll
    for (int b = 0; b < NB_B; b++) { float lo, hi; ck_unpack(AX[b], lo, hi); ax[b] = lo + hi; ck_unpack(AY[b], lo, hi); ay[b] = lo + hi; ck_unpack(AZ[b], lo, hi); az[b] = lo + hi; }
    for (int s = 1; s < NB_B; s++) {
        if (slice == s) {
#pragma unroll
            for (int b = 0; b < NB_B; b++) { red[3 * b][q] = ax[b]; red[3 * b + 1][q] = ay[b]; red[3 * b + 2][q] = az[b]; }
        }
        __syncthreads();
        if (slice == 0) {
#pragma unroll
            for (int b = 0; b < NB_B; b++) { ax[b] += red[3 * b][q]; ay[b] += red[3 * b + 1][q]; az[b] += red[3 * b + 2][q]; }
        }
        __syncthreads();
    }
    if (slice == 0) {
#pragma unroll
        for (int b = 0; b < NB_B; b++) {
            const int i = blockFirst + q * NB_B + b;
            const float4 p = pos[i];
            float4 v = vel[i];
            v.x = __fmaf_rn(ax[b], dt, v.x); v.y = __fmaf_rn(ay[b], dt, v.y); v.z = __fmaf_rn(az[b], dt, v.z);
            vel[i] = v;
            posOut[i] = make_float4(__fmaf_rn(v.x, dt, p.x), __fmaf_rn(v.y, dt, p.y), __fmaf_rn(v.z, dt, p.z), p.w);
        }
    }
}

int main() {
    const int n = 131072; std::vector<float4> h(n); unsigned s = 12345; auto rnd = [&]() { s = s * 1664525u + 1013904223u; return (s >> 8) * (1.0f / 16777216.0f) * 30.f - 15.f; };
    for (auto& p : h) { p.x = rnd(); p.y = rnd(); p.z = rnd(); p.w = 1.f; }
    float4 *d, *vel, *out; float *params, *scratch; const long long cap = 32ll << 20;   // 32 Mi floats = 128 MiB
    cudaMalloc(&d, n * 16); cudaMalloc(&vel, n * 16); cudaMalloc(&out, n * 16); cudaMalloc(&params, 64); cudaMalloc(&scratch, cap * 4);
    cudaMemcpy(d, h.data(), n * 16, cudaMemcpyHostToDevice); cudaMemset(scratch, 0, cap * 4);
    int sms, clk; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0); cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    std::vector<float4> ref(n), res(n);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    auto report = [&](const char* name, int share, float best, double maxrel) {
        double ginter = (double)share * n / best / 1e6; double frac = ginter * 1e9 * 20 / (2.0 * 128 * sms * clk * 1e3);
        printf("share %6d %-34s %7.3f ms %8.1f GInter/s  frac %.3f  maxrel %.1e  %s\n", share, name, best, ginter, frac, maxrel, cudaGetErrorString(cudaGetLastError()));
    };
    for (int share : {131072, 65536, 32768, 16384, 16384 + 32}) {
        const int first = 4096;     // a share in the middle of the array, like a device in a split
        if (share % 128 == 0) {
            cudaMemset(vel, 0, n * 16);
            float best = 1e9;
            for (int r = 0; r < 4; r++) { cudaMemset(vel, 0, n * 16); cudaEventRecord(a); nbk<<<share / 128, 128>>>(d, vel, out, n, 1e-4f, 1e-3f, share == n ? 0 : first); cudaEventRecord(b); cudaEventSynchronize(b); float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms; }
            cudaMemcpy(ref.data(), vel, n * 16, cudaMemcpyDeviceToHost);
            report("static 4x4 L128", share, best, 0);
        }
        for (int var = 0; var < 6; var++) for (int wpb : {4}) for (int L : {128}) for (int target : {16384}) {
            if ((long long)share * wpb % L) continue;
            const float hp[6] = {1e-3f, 1e-4f, (float)n, (float)cap, (float)wpb, (float)target};
            cudaMemcpy(params, hp, sizeof hp, cudaMemcpyHostToDevice);
            const int grid = (int)((long long)share * wpb / L);
            const unsigned goff = (unsigned)((share == n ? 0 : first) * wpb);
            float best = 1e9;
            for (int r = 0; r < 4; r++) { cudaMemset(vel, 0, n * 16); cudaEventRecord(a); switch (var) { case 0: nbw<2, true><<<grid, L>>>(d, vel, out, params, scratch, goff); break; case 1: nbw<4, true><<<grid, L>>>(d, vel, out, params, scratch, goff); break; case 2: nbw<2, false><<<grid, L>>>(d, vel, out, params, scratch, goff); break; case 3: nbw<4, false><<<grid, L>>>(d, vel, out, params, scratch, goff); break; case 4: nbw<1, false><<<grid, L>>>(d, vel, out, params, scratch, goff); break; default: nbw<8, false><<<grid, L>>>(d, vel, out, params, scratch, goff); } cudaEventRecord(b); cudaEventSynchronize(b); float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms; }
            cudaMemcpy(res.data(), vel, n * 16, cudaMemcpyDeviceToHost);
            double maxrel = 0; const int f0 = share == n ? 0 : first;
            if (share % 128 == 0) for (int i = f0; i < f0 + share; i++) { double dd = fabs(res[i].x - ref[i].x) / fmax(1e-5, fabs(ref[i].x)); if (dd > maxrel) maxrel = dd; }
            int untouched = 0; for (int i = 0; i < n; i++) if ((i < f0 || i >= f0 + share) && (res[i].x != 0.f || res[i].y != 0.f)) untouched++;
            int hc[2]; cudaMemcpy(hc, scratch, 8, cudaMemcpyDeviceToHost);
            char name[96]; snprintf(name, sizeof name, "queue var%d wpb%d L%d target %d%s%s", var, wpb, L, target, untouched ? " OUT-OF-RANGE-WRITES" : "", (hc[0] || hc[1]) ? " CTR-NOT-RESET" : "");
            report(name, share, best, maxrel);
        }
    }
    return 0;
}
