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
//             pos is a whole-array `read` input (every device needs all of it: the one broadcast of the path).  The
//             inverse square root is the raw SFU approximation (MUFU.RSQ); parity with the oracle is <= 1e-4 relative
//             on positions.  One work-item = one body (numberOfElementsPerWorkItem = 4 floats on vel/posOut); the
//             work-group size must be a multiple of 128.  The j loop runs over PAIRS of bodies with
//             Blackwell's packed two-lane FP32 instructions (FADD2/FMUL2/FFMA2, PTX *.f32x2): each issue slot carries
//             two interactions.  Measured on B200 (profiles/experiments micro-benchmark, round 1, per-SM cycles per
//             warp-interaction): scalar 1 body/thread 17.9, packed pairs with 2 bodies/thread 15.9; the floor is 12
//             FMA-pipe cycles (3 sub, 3 fma, 3 mul, 3 fma per interaction).  See the comment above the kernel for how
//             the work-group divides its bodies and the j range.
//   nBody3D1  the same step with scalar instructions and no restriction on the work-group size.
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

__kernel void nBody3D1(const float4* __restrict__ pos, float4* __restrict__ vel, float4* __restrict__ posOut,
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

// ---- packed two-lane FP32 helpers (sm_100+): one instruction, two independent IEEE operations ----
typedef unsigned long long ck_f32x2;
__device__ __forceinline__ ck_f32x2 ck_pack(float lo, float hi) { ck_f32x2 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void ck_unpack(ck_f32x2 v, float& lo, float& hi) { asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ ck_f32x2 ck_fma2(ck_f32x2 a, ck_f32x2 b, ck_f32x2 c) { ck_f32x2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ ck_f32x2 ck_mul2(ck_f32x2 a, ck_f32x2 b) { ck_f32x2 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ ck_f32x2 ck_sub2(ck_f32x2 a, ck_f32x2 b) { ck_f32x2 d; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }

// one body i against one PAIR of bodies j: 12 packed FMA-pipe instructions + 2 MUFU for two interactions
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

// Work decomposition of nBody3D inside a work-group of L work-items (L a multiple of 128; one work-item = one body, so the
// group owns L consecutive bodies): the group is cut into NB_B = 4 slices of L/4 threads.  Thread q of every slice carries
// the same FOUR bodies (4q .. 4q+3 of the group); slice s sweeps the s-th quarter of the j tiles, and the partial
// accelerations are combined through shared memory at the end.  Each thread thus reuses one pair of LDS.128 for 8
// interactions and keeps 4 independent dependency chains in flight, while the grid still has one thread per body.
// Measured on B200 (profiles/experiments/nb2.cu, round 1, 131072 bodies): 2 bodies per thread 8.29 ms, 4 bodies per thread 7.61 ms.
// The next tile is prefetched into registers while the current one is consumed.
#define NB_B 4          // bodies per thread = j slices per work-group
#define NB_PREFETCH 8   // float4 loads per thread per tile: NB_TILE / (L / NB_B) must not exceed this (L >= 128)

__kernel void nBody3D(const float4* __restrict__ pos, float4* __restrict__ vel, float4* __restrict__ posOut,
                      const float* __restrict__ params)
{
    // per slice: a tile of NB_TILE bodies stored as NB_TILE/2 pairs, sA[p] = (x0, x1, y0, y1), sB[p] = (z0, z1, m0, m1),
    // so one broadcast LDS.128 delivers two aligned register pairs = two packed operands
    __shared__ float4 sA[NB_B][NB_TILE / 2];
    __shared__ float4 sB[NB_B][NB_TILE / 2];
    __shared__ float red[3 * NB_B][NB_TILE];   // partial accelerations of slices 1..3 (3 per body, up to 256 threads per slice)
    const float dt = __ldg(params + 0), eps2 = __ldg(params + 1);
    const int n = (int)__ldg(params + 2);
    const int LS = blockDim.x / NB_B;                            // threads per slice
    const int slice = threadIdx.x / LS;
    const int q = threadIdx.x - slice * LS;
    const int groupFirst = (int)get_global_id(0) - threadIdx.x;  // first body of this work-group
    const int i0 = groupFirst + NB_B * q;
    float4 pb[NB_B];
    ck_f32x2 PIX[NB_B], PIY[NB_B], PIZ[NB_B], AX[NB_B], AY[NB_B], AZ[NB_B];
#pragma unroll
    for (int b = 0; b < NB_B; b++) {
        pb[b] = pos[i0 + b];
        PIX[b] = ck_pack(pb[b].x, pb[b].x);
        PIY[b] = ck_pack(pb[b].y, pb[b].y);
        PIZ[b] = ck_pack(pb[b].z, pb[b].z);
        AX[b] = ck_pack(0.0f, 0.0f);
        AY[b] = AX[b];
        AZ[b] = AX[b];
    }
    const ck_f32x2 EPS = ck_pack(eps2, eps2);

    const int tiles = (n + NB_TILE - 1) / NB_TILE;
    const int tilesPerSlice = (tiles + NB_B - 1) / NB_B;         // every slice runs the same trip count (barriers must match)
    const int firstTile = slice * tilesPerSlice;
    float* fA = reinterpret_cast<float*>(sA[slice]);
    float* fB = reinterpret_cast<float*>(sB[slice]);

    float4 nxt[NB_PREFETCH];
#pragma unroll
    for (int k = 0; k < NB_PREFETCH; k++) {
        const int t = q + k * LS;
        const int j = firstTile * NB_TILE + t;
        nxt[k] = (t < NB_TILE && firstTile < tiles && j < n) ? pos[j] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    }
    for (int it = 0; it < tilesPerSlice; it++) {
        // stage the prefetched tile (padding bodies have mass 0 and contribute nothing)
#pragma unroll
        for (int k = 0; k < NB_PREFETCH; k++) {
            const int t = q + k * LS;
            if (t < NB_TILE) {
                const int p = t >> 1, h = t & 1;
                fA[p * 4 + h] = nxt[k].x;
                fA[p * 4 + 2 + h] = nxt[k].y;
                fB[p * 4 + h] = nxt[k].z;
                fB[p * 4 + 2 + h] = nxt[k].w;
            }
        }
        __syncthreads();
        // prefetch the following tile while this one is consumed
        const int nextTile = firstTile + it + 1;
#pragma unroll
        for (int k = 0; k < NB_PREFETCH; k++) {
            const int t = q + k * LS;
            const int j = nextTile * NB_TILE + t;
            const bool live = (it + 1 < tilesPerSlice) && (nextTile < tiles) && (t < NB_TILE) && (j < n);
            nxt[k] = live ? pos[j] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        }
        const float4* tA = sA[slice];
        const float4* tB = sB[slice];
#pragma unroll 2
        for (int t = 0; t < NB_TILE / 2; t++) {
            const float4 a = tA[t], b4 = tB[t];
            const ck_f32x2 PX = ck_pack(a.x, a.y), PY = ck_pack(a.z, a.w), PZ = ck_pack(b4.x, b4.y), PM = ck_pack(b4.z, b4.w);
#pragma unroll
            for (int b = 0; b < NB_B; b++) NB_PAIR_INTERACT(PIX[b], PIY[b], PIZ[b], AX[b], AY[b], AZ[b])
        }
        __syncthreads();
    }
    float ax[NB_B], ay[NB_B], az[NB_B];
#pragma unroll
    for (int b = 0; b < NB_B; b++) {
        float lo, hi;
        ck_unpack(AX[b], lo, hi); ax[b] = lo + hi;
        ck_unpack(AY[b], lo, hi); ay[b] = lo + hi;
        ck_unpack(AZ[b], lo, hi); az[b] = lo + hi;
    }
    // slices 1..3 hand their partial sums to slice 0, one after the other (fixed order: results do not depend on timing)
    for (int s = 1; s < NB_B; s++) {
        if (slice == s) {
#pragma unroll
            for (int b = 0; b < NB_B; b++) {
                red[3 * b][q] = ax[b];
                red[3 * b + 1][q] = ay[b];
                red[3 * b + 2][q] = az[b];
            }
        }
        __syncthreads();
        if (slice == 0) {
#pragma unroll
            for (int b = 0; b < NB_B; b++) {
                ax[b] += red[3 * b][q];
                ay[b] += red[3 * b + 1][q];
                az[b] += red[3 * b + 2][q];
            }
        }
        __syncthreads();
    }
    if (slice == 0) {
#pragma unroll
        for (int b = 0; b < NB_B; b++) {
            float4 v = vel[i0 + b];
            v.x = __fmaf_rn(ax[b], dt, v.x);
            v.y = __fmaf_rn(ay[b], dt, v.y);
            v.z = __fmaf_rn(az[b], dt, v.z);
            vel[i0 + b] = v;
            posOut[i0 + b] = make_float4(__fmaf_rn(v.x, dt, pb[b].x), __fmaf_rn(v.y, dt, pb[b].y), __fmaf_rn(v.z, dt, pb[b].z), pb[b].w);
        }
    }
}

// nBody3Dq — the same step with FOUR work-items per body (global range = 4 n, numberOfElementsPerWorkItem = 1 float on
// vel/posOut, so work-item w still owns float w of those arrays and every device's share is a whole number of bodies).
// The four adjacent lanes of a body each sweep a quarter of every tile (pairs 4m + lane) and combine with two shuffles.
// It trades operand reuse for parallelism: meant for small per-device shares (an 8-way split of 131072 bodies leaves a
// device 512 warps with nBody3D — fewer than its 592 schedulers — but 2048 with this kernel).
__kernel void nBody3Dq(const float4* __restrict__ pos, float4* __restrict__ vel, float4* __restrict__ posOut,
                       const float* __restrict__ params)
{
    __shared__ float4 sA[NB_TILE / 2];
    __shared__ float4 sB[NB_TILE / 2];
    float* fA = reinterpret_cast<float*>(sA);
    float* fB = reinterpret_cast<float*>(sB);
    const float dt = __ldg(params + 0), eps2 = __ldg(params + 1);
    const int n = (int)__ldg(params + 2);
    const unsigned int w = (unsigned int)get_global_id(0);
    const int i = (int)(w >> 2), lane4 = (int)(w & 3u);
    const float4 pi = pos[i];
    const ck_f32x2 PIX = ck_pack(pi.x, pi.x), PIY = ck_pack(pi.y, pi.y), PIZ = ck_pack(pi.z, pi.z), EPS = ck_pack(eps2, eps2);
    ck_f32x2 AX = ck_pack(0.0f, 0.0f), AY = AX, AZ = AX;
    for (int base = 0; base < n; base += NB_TILE) {
        for (int t = threadIdx.x; t < NB_TILE; t += blockDim.x) {
            const int j = base + t;
            const float4 pj = (j < n) ? pos[j] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            const int p = t >> 1, h = t & 1;
            fA[p * 4 + h] = pj.x;
            fA[p * 4 + 2 + h] = pj.y;
            fB[p * 4 + h] = pj.z;
            fB[p * 4 + 2 + h] = pj.w;
        }
        __syncthreads();
#pragma unroll 4
        for (int m = 0; m < NB_TILE / 8; m++) {
            const float4 a = sA[4 * m + lane4], b4 = sB[4 * m + lane4];   // a warp reads 64 contiguous bytes: one wavefront
            const ck_f32x2 PX = ck_pack(a.x, a.y), PY = ck_pack(a.z, a.w), PZ = ck_pack(b4.x, b4.y), PM = ck_pack(b4.z, b4.w);
            NB_PAIR_INTERACT(PIX, PIY, PIZ, AX, AY, AZ)
        }
        __syncthreads();
    }
    float lo, hi;
    ck_unpack(AX, lo, hi); float ax = lo + hi;
    ck_unpack(AY, lo, hi); float ay = lo + hi;
    ck_unpack(AZ, lo, hi); float az = lo + hi;
    ax += __shfl_xor_sync(0xffffffffu, ax, 1); ay += __shfl_xor_sync(0xffffffffu, ay, 1); az += __shfl_xor_sync(0xffffffffu, az, 1);
    ax += __shfl_xor_sync(0xffffffffu, ax, 2); ay += __shfl_xor_sync(0xffffffffu, ay, 2); az += __shfl_xor_sync(0xffffffffu, az, 2);
    if (lane4 == 0) {
        float4 v = vel[i];
        v.x = __fmaf_rn(ax, dt, v.x);
        v.y = __fmaf_rn(ay, dt, v.y);
        v.z = __fmaf_rn(az, dt, v.z);
        vel[i] = v;
        posOut[i] = make_float4(__fmaf_rn(v.x, dt, pi.x), __fmaf_rn(v.y, dt, pi.y), __fmaf_rn(v.z, dt, pi.z), pi.w);
    }
}

// nBody3Dw — the same step scheduled through a device-side work queue, so that the time of a launch does not depend on how
// the launch geometry happens to divide over 148 SMs.  With one equal slab of work per work-group (nBody3D) a device's share
// of 131072 / D bodies gives 1024 / D work-groups for 592 resident slots: 1.73, 0.86, 0.43 ... waves, i.e. 13-57 % of the
// SM-time idle in the tail (ncu on nBody3D, round 1: 13.3 of 16 warps active on average).  Here the work is cut into UNITS
//     unit (g, c) = bodies [128 g, 128 g + 128) of the share  x  j tiles [c T, c T + T),   T = NBW_TILES_PER_WARP * warps,
// and every work-group pulls units from a counter until none are left (work-groups scheduled after that exit at once).
// A unit is computed like a work-group of nBody3D (lane q carries bodies 4q..4q+3, warp w sweeps its share of the tiles,
// packed pair arithmetic), its 128 x 3 partial accelerations go to `scratch`, and the work-group that completes the LAST
// unit of a body group adds the partials in chunk order and integrates the group.  Sums therefore do not depend on timing,
// nor on how the bodies are split over devices (T depends on n and the work-group size only) — unless scratch is too small
// for that chunking, in which case chunks are widened to fit.
//   arguments   pos (whole `read`), vel, posOut (sliced; numberOfElementsPerWorkItem = 4 / workItemsPerBody floats),
//               params = { dt, eps2, n, workItemsPerBody (1, 2 or 4), scratch capacity in units of 1024 floats },
//               scratch: float array with read = write = false (device-resident, zeroed when the buffer is created).
//   geometry    global range = workItemsPerBody * bodies; work-group size a multiple of 32 up to 256 (128 measured best).
//               More work-items per body only means more resident work-groups to pull units: with 4 an 8-way split of
//               131072 bodies still launches 512 work-groups per device.
//   scratch     [0] next unit, [1] work-groups that left, [4 .. 4+G) units done per body group, then G*C*384 floats of
//               partials (G = groups of the share, C = chunks); 131072 bodies on one device: C = 64, 100.7 MB.  The kernel
//               leaves all counters at zero again.  One launch at a time per scratch array: do not use it from pipelined
//               computes (blobs > 1) or from two computeIds at once.
// Measured on B200, round 1, GInter/s per device for a share of 131072 / 65536 / 32768 / 16384 bodies:
//   nBody3D (stand-alone harness) 2200 / 1791 / 1578 / 882;  a prototype of this kernel with coarser chunks in the same
//   harness 2591 / 2580 / 2536 / 2465;  this kernel through bench.py 2570 (69.0 % of the FP32 peak) / 2556 on 1 / 2 devices.
// What keeps it from the 12-cycle-per-interaction floor of the FMA pipe (83 % of peak): the three accumulating FFMA2 of a
// pair read three distinct 64-bit register pairs each, which costs 3 cycles instead of 2 on Blackwell's two register banks
// (profiles/experiments/f2.cu), i.e. 13.97 cycles per interaction = 71.6 %; the rest is the per-unit hand-over.
#define NBW_GROUP 128            // bodies per unit: 32 lanes x NB_B
#define NBW_MAXW 8               // warps per work-group at most
#define NBW_HDR 4                // ints before the per-group counters
#ifndef NBW_TILES_PER_WARP
#define NBW_TILES_PER_WARP 4     // j tiles a warp sweeps per unit (round 2: 2 -> 69.2 %, 4 -> 69.5 %, 8 -> 69.6 % of the FP32 peak on one B200)
#endif

// The step itself, shared by nBody3Dw and nBody3Dwp (below).  With NBW_PEERS the integrating work-group also stores each
// new position into the posOut buffers of the other devices (peers[0] = count, peers[1..] = device pointers, NVLink peer
// stores), so that after the launch EVERY device holds the complete new positions and the next step needs no host.
template <bool NBW_PEERS>
__device__ __forceinline__ void ck_nbw_step(const unsigned int __ck_goff, const float4* __restrict__ pos, float4* __restrict__ vel,
                                            float4* __restrict__ posOut, const float* __restrict__ params,
                                            float* __restrict__ scratch, const unsigned long long* __restrict__ peers)
{
    __shared__ float4 sA[NBW_MAXW][NB_TILE / 2];
    __shared__ float4 sB[NBW_MAXW][NB_TILE / 2];
    __shared__ float red[NBW_MAXW - 1][3 * NB_B][32];
    __shared__ int sNext, sLast;
    const float dt = __ldg(params + 0), eps2 = __ldg(params + 1);
    const int n = (int)__ldg(params + 2);
    const int wpb = (int)__ldg(params + 3);
    const long long cap = (long long)__ldg(params + 4) * 1024;
    const int S = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int firstBody = (int)(get_global_offset(0) / wpb);
    const int share = (int)((long long)gridDim.x * blockDim.x / wpb);
    const int lastBody = firstBody + share - 1;
    const int G = (share + NBW_GROUP - 1) / NBW_GROUP;
    const int tiles = (n + NB_TILE - 1) / NB_TILE;
    int* ctr = reinterpret_cast<int*>(scratch);
    int* done = ctr + NBW_HDR;
    const long long P0 = ((long long)(NBW_HDR + G) + 127) / 128 * 128;
    // chunking: T tiles per unit; widened (in whole multiples of the warp count) only if scratch cannot hold G*C partial sets
    int T = NBW_TILES_PER_WARP * S;
    int C = (tiles + T - 1) / T;
    {
        const long long cfit = (cap - P0) / ((long long)G * 3 * NBW_GROUP);
        if (cfit < 1 || S < 1 || S > NBW_MAXW || share < NB_B) {
            if (blockIdx.x == 0 && threadIdx.x == 0)
                printf("nBody3Dw: scratch holds %lld floats, needs at least %lld; work-group size %d must be 32..256\n", cap,
                       P0 + (long long)G * 3 * NBW_GROUP, (int)blockDim.x);
            return;
        }
        if (C > cfit) {
            T = (int)(((tiles + cfit - 1) / cfit + S - 1) / S) * S;
            C = (tiles + T - 1) / T;
        }
    }
    const int totalUnits = G * C;
    float* part = scratch + P0;
    const ck_f32x2 EPS = ck_pack(eps2, eps2);
    float* fA = reinterpret_cast<float*>(sA[warp]);
    float* fB = reinterpret_cast<float*>(sB[warp]);
    const float4* tA = sA[warp];
    const float4* tB = sB[warp];

    // The work-group always knows its NEXT unit as well as the current one (it draws two tickets at the start and one
    // per finished unit), so the first j tile of the next unit is prefetched while the last tile of this one is consumed.
    float4 nxt[NB_PREFETCH];
    auto fetchTile = [&](int tile, bool live) {
#pragma unroll
        for (int k = 0; k < NB_PREFETCH; k++) {
            const int j = tile * NB_TILE + lane + 32 * k;
            nxt[k] = (live && j < n) ? pos[j] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);   // padding bodies have mass 0
        }
    };
    const int per = T / S;
    auto warpTiles = [&](int unit, int& first, int& end) {   // this warp's tiles of the unit's chunk
        const int c = unit % C;
        const int chunkEnd = min(c * T + T, tiles);
        first = min(c * T + warp * per, chunkEnd);
        end = min(first + per, chunkEnd);
    };
    if (threadIdx.x == 0) {
        sNext = atomicAdd(&ctr[0], 1);
        sLast = atomicAdd(&ctr[0], 1);
    }
    __syncthreads();
    int u = sNext, un = sLast;
    __syncthreads();
    int t0 = 0, t1 = 0, gLoaded = -1;
    if (u < totalUnits) warpTiles(u, t0, t1);
    fetchTile(t0, t0 < t1);
    ck_f32x2 PIX[NB_B], PIY[NB_B], PIZ[NB_B], AX[NB_B], AY[NB_B], AZ[NB_B];
    while (u < totalUnits) {
        const int g = u / C;
        if (g != gLoaded) {   // consecutive units mostly stay in one body group
            int i0 = firstBody + g * NBW_GROUP + NB_B * lane;
            if (i0 > lastBody) i0 = lastBody - (NB_B - 1);   // lanes past a ragged last group recompute live bodies; nothing of theirs is used
#pragma unroll
            for (int b = 0; b < NB_B; b++) {
                const float4 p = pos[i0 + b];
                PIX[b] = ck_pack(p.x, p.x);
                PIY[b] = ck_pack(p.y, p.y);
                PIZ[b] = ck_pack(p.z, p.z);
            }
            gLoaded = g;
        }
#pragma unroll
        for (int b = 0; b < NB_B; b++) {
            AX[b] = ck_pack(0.0f, 0.0f);
            AY[b] = AX[b];
            AZ[b] = AX[b];
        }
        int n0 = 0, n1 = 0;
        if (un < totalUnits) warpTiles(un, n0, n1);
        if (t0 >= t1) fetchTile(n0, n0 < n1);   // a short last chunk left this warp nothing: go straight to the next unit's tile
        for (int tile = t0; tile < t1; tile++) {
#pragma unroll
            for (int k = 0; k < NB_PREFETCH; k++) {
                const int t = lane + 32 * k;
                const int p = t >> 1, h = t & 1;
                fA[p * 4 + h] = nxt[k].x;
                fA[p * 4 + 2 + h] = nxt[k].y;
                fB[p * 4 + h] = nxt[k].z;
                fB[p * 4 + 2 + h] = nxt[k].w;
            }
            __syncwarp();
            if (tile + 1 < t1) fetchTile(tile + 1, true);
            else fetchTile(n0, n0 < n1);
            // unroll 4: with 2 the loop-carried accumulators cost 16 register moves per iteration (SASS, round 1)
#pragma unroll 4
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
            for (int s = 0; s < S - 1; s++) {   // fixed order: warp 1, 2, ...
#pragma unroll
                for (int k = 0; k < 3 * NB_B; k++) acc[k] += red[s][k][lane];
            }
            float* dst = part + (long long)u * (3 * NBW_GROUP);
#pragma unroll
            for (int b = 0; b < NB_B; b++) {
                __stcg(dst + 0 * NBW_GROUP + b * 32 + lane, acc[3 * b]);
                __stcg(dst + 1 * NBW_GROUP + b * 32 + lane, acc[3 * b + 1]);
                __stcg(dst + 2 * NBW_GROUP + b * 32 + lane, acc[3 * b + 2]);
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
            // every unit of group g has published its partials (fence before the counter, fence after it here)
            __threadfence();
            for (int t = threadIdx.x; t < NBW_GROUP; t += blockDim.x) {
                const int body = firstBody + g * NBW_GROUP + NB_B * (t & 31) + (t >> 5);
                if (body <= lastBody) {
                    float ax = 0.0f, ay = 0.0f, az = 0.0f;
                    const float* src = part + (long long)g * C * (3 * NBW_GROUP) + t;
#pragma unroll 8
                    for (int cc = 0; cc < C; cc++) {
                        ax += __ldcg(src + 0 * NBW_GROUP);
                        ay += __ldcg(src + 1 * NBW_GROUP);
                        az += __ldcg(src + 2 * NBW_GROUP);
                        src += 3 * NBW_GROUP;
                    }
                    const float4 p = pos[body];
                    float4 v = vel[body];
                    v.x = __fmaf_rn(ax, dt, v.x);
                    v.y = __fmaf_rn(ay, dt, v.y);
                    v.z = __fmaf_rn(az, dt, v.z);
                    vel[body] = v;
                    const float4 np = make_float4(__fmaf_rn(v.x, dt, p.x), __fmaf_rn(v.y, dt, p.y), __fmaf_rn(v.z, dt, p.z), p.w);
                    posOut[body] = np;
                    if (NBW_PEERS) {
                        // 16-byte stores over NVLink, two per body and peer: the position for the next step, and the velocity
                        // so that the body can change owner when the balancer moves the split
                        const int count = (int)peers[0];
                        for (int d = 1; d <= count; d++) {
                            float4* remotePos = reinterpret_cast<float4*>(peers[d]);
                            float4* remoteVel = reinterpret_cast<float4*>(peers[count + d]);
                            if (remotePos != posOut) {
                                remotePos[body] = np;
                                remoteVel[body] = v;
                            }
                        }
                    }
                }
            }
            if (threadIdx.x == 0) done[g] = 0;
        }
        __syncthreads();
        u = un;
        un = unext;
        t0 = n0;
        t1 = n1;
    }
    // the last work-group to leave re-arms the queue for the next launch (every work-group draws its tickets before it
    // leaves, and leaves once)
    if (threadIdx.x == 0) {
        if (atomicAdd(&ctr[1], 1) == (int)gridDim.x - 1) {
            ctr[0] = 0;
            ctr[1] = 0;
        }
    }
}

__kernel void nBody3Dw(const float4* __restrict__ pos, float4* __restrict__ vel, float4* __restrict__ posOut,
                       const float* __restrict__ params, float* __restrict__ scratch)
{
    ck_nbw_step<false>((unsigned int)get_global_offset(0), pos, vel, posOut, params, scratch, nullptr);
}

// nBody3Dwp — the step fused with the exchange that follows it when the bodies are split over several devices and stay
// there: compute (range split) + all-gather of the new positions in ONE kernel, over peer memory.  In the reference's model
// every step round-trips through the host (download the slices, upload the whole array to every device: Worker.cs:833-834);
// with `peers` = { D, posOut on device 0..D-1, vel on device 0..D-1 } (device addresses; ClNumberCruncher.peerTable(posOut,
// vel) in api.py fills it) the integrating work-groups write each new position and velocity to every device, and the next
// step runs with read/write flags off.  Every device then holds complete arrays, so the balancer may move the split freely.
// The host-side finish of all devices at the end of compute() is the only synchronisation needed between steps.
__kernel void nBody3Dwp(const float4* __restrict__ pos, float4* __restrict__ vel, float4* __restrict__ posOut,
                        const float* __restrict__ params, float* __restrict__ scratch,
                        const unsigned long long* __restrict__ peers)
{
    ck_nbw_step<true>((unsigned int)get_global_offset(0), pos, vel, posOut, params, scratch, peers);
}
