// oracle.cpp — CPU restatement of the kernels on the ClNumberCruncher hot path.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under cekirdekler_b200/ links, loads or calls this file; only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do, as the checker and the
// timed CPU baseline.  The product path has no CPU execution route.
//
// What is restated and from where (paths relative to /root/reference/Cekirdekler/Cekirdekler/):
//   oracle_copy         the `test` copy kernels                      Tester.cs:35-38 (and the 6 other element types)
//   oracle_vector_add   `vectorAdd`  c[i]=a[i]+b[i]                  Tester.cs:7820-7824
//   oracle_nbody2d      the host n-body loop = the `nBody` kernel    Tester.cs:7702-7716 / :7727-7742 (quirk Q3 kept: fy uses dx)
//   oracle_wave_equation  the CPU branch of the Unity sample         /root/reference/Kamera.cs:213-218
// Kernels that BASELINE.json names but the reference does not contain have NO reference source and NO
// reference vectors: their arithmetic is DEFINED here and in cekirdekler_b200/csrc/kernels/*.cu, operation by operation,
// so both sides can be bit-exact.  PARITY UNPINNED for: oracle_mandelbrot, oracle_saxpy, oracle_nbody3d_step,
// oracle_blur5x5, oracle_sobel, oracle_histogram.
//
// Build: g++ -O3 -march=x86-64-v3 -ffp-contract=off -fopenmp (see cekirdekler_b200/_build.py).  With contraction off,
// the only fused multiply-adds are the explicit fmaf() calls, which mirror the explicit __fmaf_rn() of the CUDA kernels.
#include <cmath>
#include <cstdint>
#include <cstring>

#ifdef _OPENMP
#include <omp.h>
#endif

extern "C" {

int oracle_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

void oracle_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

// ---------------------------------------------------------------- K6: copy (Tester.cs:35-38)
// __kernel void test(__global T* data, __global T* data2){ int id=get_global_id(0); data2[id]=data[id]; }
void oracle_copy(const void* data, void* data2, long long offset, long long count, int elemSize) {
    memcpy((char*)data2 + offset * elemSize, (const char*)data + offset * elemSize, (size_t)(count * elemSize));
}

// ---------------------------------------------------------------- K2: vectorAdd (Tester.cs:7820-7824)
void oracle_vector_add(const float* a, const float* b, float* c, long long offset, long long count) {
#pragma omp parallel for schedule(static)
    for (long long i = offset; i < offset + count; i++) c[i] = a[i] + b[i];
}

// ---------------------------------------------------------------- K3: SAXPY y = alpha*x + y (defined here; no reference source)
// Two roundings (multiply, then add): the CUDA kernel uses __fmul_rn/__fadd_rn, i.e. contraction disabled.
void oracle_saxpy(const float* alpha, const float* x, float* y, long long offset, long long count) {
    const float a = alpha[0];
#pragma omp parallel for schedule(static)
    for (long long i = offset; i < offset + count; i++) {
        float p = a * x[i];
        y[i] = p + y[i];
    }
}

// ---------------------------------------------------------------- K4 (reference-faithful 2-D): Tester.cs:7702-7716
// One work-item per body i over ALL n bodies j (self-interaction included).  Plain float ops in source order,
// no contraction; fy accumulates dx exactly as the reference does (quirk Q3).
void oracle_nbody2d(const float* x, const float* y, float* fx, float* fy, int n, long long offset, long long count) {
#pragma omp parallel for schedule(static)
    for (long long i = offset; i < offset + count; i++) {
        float fx0 = 0, fy0 = 0;
        for (int j = 0; j < n; j++) {
            float dx = x[i] - x[j];
            float dy = y[i] - y[j];
            float r = sqrtf(dx * dx + dy * dy + 0.0001f);
            fx0 += (dx * (1.0f / (r * r * r)));
            fy0 += (dx * (1.0f / (r * r * r)));
        }
        fx[i] = fx0;
        fy[i] = fy0;
    }
}

// ---------------------------------------------------------------- K1: Mandelbrot escape time (defined here)
// view = { x0, y0, dx, dy, (float)W, (float)maxIter, 0, 0 }.  Pixel p: px = p % W, py = p / W,
//   cx = fmaf(px, dx, x0), cy = fmaf(py, dy, y0), z0 = 0.
// out[p] = the first k in [0, maxIter) with fmaf(x_k, x_k, y_k*y_k) > 4, else maxIter, where
//   x_{k+1} = fmaf(x_k, x_k, fmaf(-y_k, y_k, cx)),   y_{k+1} = fmaf(2, x_k*y_k, cy).
// 4 rounded operations per iteration (+2 for the escape test); by convention the work is counted as 8 FLOP per executed iteration
// (3 mul + 5 add of the textbook form).
static inline int mandel_pixel(float cx, float cy, int maxIter) {
    float x = 0.0f, y = 0.0f;
    int k = 0;
    while (k < maxIter) {
        float m = fmaf(x, x, y * y);
        if (m > 4.0f) break;
        float t = fmaf(-y, y, cx);
        float w = x * y;
        y = fmaf(2.0f, w, cy);
        x = fmaf(x, x, t);
        k++;
    }
    return k;
}

void oracle_mandelbrot(const float* view, int32_t* out, long long offset, long long count) {
    const float x0 = view[0], y0 = view[1], dx = view[2], dy = view[3];
    const long long W = (long long)view[4];
    const int maxIter = (int)view[5];
#pragma omp parallel for schedule(dynamic, 1024)
    for (long long p = offset; p < offset + count; p++) {
        const float cx = fmaf((float)(p % W), dx, x0);
        const float cy = fmaf((float)(p / W), dy, y0);
        out[p] = mandel_pixel(cx, cy, maxIter);
    }
}

// strided sample for the bounded CPU baseline: pixels p = offset + s*stride, s in [0,count); returns Σ iterations
long long oracle_mandelbrot_sample(const float* view, long long offset, long long stride, long long count, int32_t* outOrNull) {
    const float x0 = view[0], y0 = view[1], dx = view[2], dy = view[3];
    const long long W = (long long)view[4];
    const int maxIter = (int)view[5];
    long long total = 0;
#pragma omp parallel for schedule(dynamic, 256) reduction(+ : total)
    for (long long s = 0; s < count; s++) {
        const long long p = offset + s * stride;
        const float cx = fmaf((float)(p % W), dx, x0);
        const float cy = fmaf((float)(p / W), dy, y0);
        const int k = mandel_pixel(cx, cy, maxIter);
        if (outOrNull) outOrNull[s] = k;
        total += k;
    }
    return total;
}

// ---------------------------------------------------------------- K4 (BASELINE float4, 3-D): one leap-frog step (defined here)
// pos[i] = (x, y, z, mass), vel[i] = (vx, vy, vz, unused); params = { dt, eps2 }.
//   a_i = Σ_j m_j * d / (|d|^2 + eps2)^(3/2),  d = p_j - p_i   (j over all n bodies, in index order, self included: d=0 adds 0)
//   v_i += a_i * dt ;  posOut_i = p_i + v_i * dt ; mass is carried over.
// Per interaction: d (3 sub), r2 = fma chain (3), inv = 1/sqrt(r2), inv3 = inv*inv*inv (2), s = m*inv3 (1), acc fma (3).
void oracle_nbody3d_step(const float* pos, float* vel, float* posOut, const float* params, int n, long long offset, long long count) {
    const float dt = params[0], eps2 = params[1];
#pragma omp parallel for schedule(static)
    for (long long i = offset; i < offset + count; i++) {
        const float xi = pos[4 * i], yi = pos[4 * i + 1], zi = pos[4 * i + 2];
        float ax = 0, ay = 0, az = 0;
        for (int j = 0; j < n; j++) {
            const float dx = pos[4 * j] - xi, dy = pos[4 * j + 1] - yi, dz = pos[4 * j + 2] - zi;
            const float r2 = fmaf(dz, dz, fmaf(dy, dy, fmaf(dx, dx, eps2)));
            const float inv = 1.0f / sqrtf(r2);
            const float s = pos[4 * j + 3] * (inv * inv * inv);
            ax = fmaf(dx, s, ax);
            ay = fmaf(dy, s, ay);
            az = fmaf(dz, s, az);
        }
        const float vx = fmaf(ax, dt, vel[4 * i]), vy = fmaf(ay, dt, vel[4 * i + 1]), vz = fmaf(az, dt, vel[4 * i + 2]);
        vel[4 * i] = vx;
        vel[4 * i + 1] = vy;
        vel[4 * i + 2] = vz;
        posOut[4 * i] = fmaf(vx, dt, xi);
        posOut[4 * i + 1] = fmaf(vy, dt, yi);
        posOut[4 * i + 2] = fmaf(vz, dt, zi);
        posOut[4 * i + 3] = pos[4 * i + 3];
    }
}

// ---------------------------------------------------------------- waveEquation (Kamera.cs:213-218, the reference's own CPU loop)
// vertices[i] = verticesBase[i] + 0.02f * normals[i] * ctr * (float)sin(40 t + 100 sqrt(dx^2 + dy^2)),  dx/dy measured from (x, y);
// arguments = { ctr, t, x, y, n } as the kernel receives them (Kamera.cs:257-262).  Vector3 arrays are plain xyz float triples.
void oracle_wave_equation(const float* xyz, const float* xyzn, float* xyzo, const float* arguments, long long offset, long long count) {
    const float ctr = arguments[0], t = arguments[1], x = arguments[2], y = arguments[3];
    const long long n = (long long)arguments[4];
#pragma omp parallel for schedule(static)
    for (long long i = offset; i < offset + count; i++) {
        if (i >= n) continue;
        const float dx = xyz[3 * i] - x, dy = xyz[3 * i + 1] - y;
        const float wave = 0.02f * ctr * (float)sin(40.0f * t + 100.0f * sqrt((double)(dx * dx + dy * dy)));
        xyzo[3 * i] = xyz[3 * i] + xyzn[3 * i] * wave;
        xyzo[3 * i + 1] = xyz[3 * i + 1] + xyzn[3 * i + 1] * wave;
        xyzo[3 * i + 2] = xyz[3 * i + 2] + xyzn[3 * i + 2] * wave;
    }
}

// ---------------------------------------------------------------- K5a: 5x5 binomial blur, clamp-to-edge (defined here)
// dims = { W, H }.  Separable [1 4 6 4 1]/16 in both directions:
//   h(x,y) = fmaf(6, in(x,y), fmaf(4, in(x-1,y)+in(x+1,y), in(x-2,y)+in(x+2,y)))
//   out(x,y) = (1/256) * fmaf(6, h(x,y), fmaf(4, h(x,y-1)+h(x,y+1), h(x,y-2)+h(x,y+2)))
static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
static inline float blur_h(const float* in, int W, int x, int y) {
    const float* r = in + (long long)y * W;
    const float c = r[x];
    const float s1 = r[clampi(x - 1, 0, W - 1)] + r[clampi(x + 1, 0, W - 1)];
    const float s2 = r[clampi(x - 2, 0, W - 1)] + r[clampi(x + 2, 0, W - 1)];
    return fmaf(6.0f, c, fmaf(4.0f, s1, s2));
}
void oracle_blur5x5(const float* in, float* out, const int32_t* dims, long long offset, long long count) {
    const int W = dims[0], H = dims[1];
#pragma omp parallel for schedule(static)
    for (long long p = offset; p < offset + count; p++) {
        const int x = (int)(p % W), y = (int)(p / W);
        const float c = blur_h(in, W, x, y);
        const float s1 = blur_h(in, W, x, clampi(y - 1, 0, H - 1)) + blur_h(in, W, x, clampi(y + 1, 0, H - 1));
        const float s2 = blur_h(in, W, x, clampi(y - 2, 0, H - 1)) + blur_h(in, W, x, clampi(y + 2, 0, H - 1));
        out[p] = 0.00390625f * fmaf(6.0f, c, fmaf(4.0f, s1, s2));
    }
}

// ---------------------------------------------------------------- K5b: 3x3 Sobel magnitude, clamp-to-edge (defined here)
//   gx = (p(+1,-1) + 2 p(+1,0) + p(+1,+1)) - (p(-1,-1) + 2 p(-1,0) + p(-1,+1)),  gy likewise on rows;
//   out = sqrtf(fmaf(gx, gx, gy*gy));  column/row sums are fmaf(2, mid, a + b).
void oracle_sobel(const float* in, float* out, const int32_t* dims, long long offset, long long count) {
    const int W = dims[0], H = dims[1];
#pragma omp parallel for schedule(static)
    for (long long p = offset; p < offset + count; p++) {
        const int x = (int)(p % W), y = (int)(p / W);
        const int xm = clampi(x - 1, 0, W - 1), xp = clampi(x + 1, 0, W - 1);
        const int ym = clampi(y - 1, 0, H - 1), yp = clampi(y + 1, 0, H - 1);
        const float* rm = in + (long long)ym * W;
        const float* r0 = in + (long long)y * W;
        const float* rp = in + (long long)yp * W;
        const float right = fmaf(2.0f, r0[xp], rm[xp] + rp[xp]);
        const float left = fmaf(2.0f, r0[xm], rm[xm] + rp[xm]);
        const float bottom = fmaf(2.0f, rp[x], rp[xm] + rp[xp]);
        const float top = fmaf(2.0f, rm[x], rm[xm] + rm[xp]);
        const float gx = right - left, gy = bottom - top;
        out[p] = sqrtf(fmaf(gx, gx, gy * gy));
    }
}

// ---------------------------------------------------------------- K5c: 256-bin histogram (defined here)
// bin = (int)min(max(v * 255.0f, 0), 255)  (clamped in float first, so no int overflow; NaN → 0); hist is overwritten (not accumulated) with the counts of [offset, offset+count).
void oracle_histogram(const float* in, uint32_t* hist, long long offset, long long count) {
    uint32_t local[256];
    memset(local, 0, sizeof local);
    for (long long p = offset; p < offset + count; p++) {
        const float v = fminf(fmaxf(in[p] * 255.0f, 0.0f), 255.0f);
        local[(int)v]++;
    }
    memcpy(hist, local, sizeof local);
}

}  // extern "C"
