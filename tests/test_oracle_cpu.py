"""CPU-only checks of the oracle itself: against the values the reference fixes (copy = identity, vectorAdd, the 2-D
n-body host loop restated independently in float64 numpy), against closed-form facts, and against the committed golden
fixtures.  No GPU needed."""
import hashlib
import json
from pathlib import Path

import numpy as np

from oracle import pyoracle as po

GOLDEN = json.loads((Path(__file__).resolve().parent / "golden" / "kernel_vectors.json").read_text())


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_copy_is_identity_on_the_requested_range_only():
    src = np.arange(1000, dtype=np.int64) * 3 - 7
    dst = np.zeros(1000, np.int64)
    po.copy(src, dst, 100, 300)
    assert (dst[100:400] == src[100:400]).all() and (dst[:100] == 0).all() and (dst[400:] == 0).all()


def test_vector_add_reference_check_value():
    # Tester.cs:7812-7833: a = 3, b = i → c[700] = 703
    n = 1 << 16
    a, b, c = np.full(n, 3, np.float32), np.arange(n, dtype=np.float32), np.full(n, 105, np.float32)
    po.vector_add(a, b, c)
    assert c[700] == 703.0 and np.array_equal(c, a + b)


def test_saxpy_is_two_roundings_not_an_fma():
    alpha = np.array([1.0 + 2.0 ** -12], np.float32)
    x = np.array([1.0 + 2.0 ** -12], np.float32)
    y = np.array([-1.0], np.float32)
    po.saxpy(alpha, x, y)
    prod = np.float32(alpha[0] * x[0])                 # rounded product
    assert y[0] == np.float32(prod + np.float32(-1.0))
    fused = np.float32(np.float64(alpha[0]) * np.float64(x[0]) - 1.0)
    assert y[0] != fused                                # an FMA would keep the low product bits


def test_nbody2d_follows_the_reference_host_loop():
    n = 256
    x, y = po.uniform(1, n, -15, 15), po.uniform(2, n, -15, 15)
    fx, fy = np.zeros(n, np.float32), np.zeros(n, np.float32)
    po.nbody2d(x, y, fx, fy)
    # independent float64 restatement of Tester.cs:7702-7716 (fy accumulates dx: quirk Q3)
    dx = x[:, None].astype(np.float64) - x[None, :]
    dy = y[:, None].astype(np.float64) - y[None, :]
    r = np.sqrt(dx * dx + dy * dy + 0.0001)
    f = (dx / r ** 3).sum(axis=1)
    assert np.allclose(fx, f, rtol=2e-3, atol=2e-3)
    assert np.array_equal(fx, fy)                       # the quirk: both components hold the x force
    g = GOLDEN["nbody2d_1024"]
    n = 1024
    x, y = po.uniform(g["seed_x"], n, -15, 15), po.uniform(g["seed_y"], n, -15, 15)
    fx, fy = np.zeros(n, np.float32), np.zeros(n, np.float32)
    po.nbody2d(x, y, fx, fy)
    assert sha(fx) == g["sha256_fx"]


def test_mandelbrot_known_points():
    # c = 0 never escapes; c = 1: z = 0,1,2,5 → |z_3|^2 = 25 > 4 at k = 3; c = -2 stays at |z| = 2 (not > 4)
    def one(cx, cy, it):
        v = np.array([cx, cy, 0, 0, 1, it, 0, 0], np.float32)
        return int(po.mandelbrot(v, 1)[0])
    assert one(0, 0, 100) == 100
    assert one(1, 0, 100) == 3
    assert one(-2, 0, 500) == 500
    assert one(2.5, 0, 100) == 1
    assert one(0.25, 0, 5000) == 5000     # cusp of the cardioid: converges to 0.5


def test_mandelbrot_golden_and_symmetry():
    g = GOLDEN["mandelbrot_c1"]
    v = po.mandelbrot_view(1024, 1024, 256)
    assert v.tolist() == g["view"]
    m = po.mandelbrot(v, 1024 * 1024)
    assert sha(m) == g["sha256"] and int(m.sum()) == g["sum"]
    img = m.reshape(1024, 1024)
    assert np.array_equal(img[1:512], img[1023:512:-1])
    # ranged evaluation equals the same slice of the full evaluation (what a device's share computes)
    part = np.zeros(1024 * 1024, np.int32)
    po.mandelbrot(v, 1024 * 1024, 300 * 1024, 5 * 1024, part)
    assert np.array_equal(part[300 * 1024:305 * 1024], m[300 * 1024:305 * 1024])
    g3 = GOLDEN["mandelbrot_c3_sample"]
    s = np.zeros(4096, np.int32)
    po.mandelbrot_sample(np.array(g3["view"], np.float32), g3["offset"], g3["stride"], 4096, s)
    full = np.zeros(g3["count"], np.int32)
    # the first 4096 entries of the committed sample must reproduce (cheap prefix of the 65536-point fixture)
    po.mandelbrot_sample(np.array(g3["view"], np.float32), g3["offset"], g3["stride"], g3["count"], full)
    assert np.array_equal(full[:4096], s) and sha(full) == g3["sha256"]


def test_nbody3d_momentum_and_golden():
    n = 512
    pos = np.zeros((n, 4), np.float32)
    for c in range(3):
        pos[:, c] = po.uniform(40 + c, n, -15, 15)
    pos[:, 3] = 1.0
    vel, out = np.zeros(4 * n, np.float32), np.zeros(4 * n, np.float32)
    po.nbody3d_step(pos.reshape(-1), vel, out, np.array([1e-3, 1e-4, n, 0], np.float32))
    # equal masses, pairwise antisymmetric forces: total momentum stays ~0
    assert np.abs(vel.reshape(n, 4)[:, :3].sum(axis=0)).max() < 1e-3
    assert np.array_equal(out.reshape(n, 4)[:, 3], pos[:, 3])


def test_image_oracles():
    W, H = 256, 128
    g = GOLDEN["image_256x128"]
    img = po.uniform(g["seed"], W * H, 0, 1)
    dims = np.array([W, H], np.int32)
    b, s, hist = np.zeros_like(img), np.zeros_like(img), np.zeros(256, np.uint32)
    po.blur5x5(img, b, dims)
    po.sobel(b, s, dims)
    po.histogram(s, hist)
    assert sha(b) == g["sha256_blur"] and sha(s) == g["sha256_sobel"] and hist.tolist() == g["hist"]
    assert int(hist.sum()) == W * H
    flat = np.full(W * H, 0.5, np.float32)
    po.blur5x5(flat, b, dims)
    assert np.array_equal(b, flat)                      # weights sum to 256/256 exactly
    po.sobel(flat, s, dims)
    assert not s.any()                                  # no gradient in a constant image
    ramp = np.tile(np.arange(W, dtype=np.float32) / W, H)
    po.sobel(ramp, s, dims)
    assert np.allclose(s.reshape(H, W)[5, 5:-5], 8.0 / W)   # |gx| = 8 * slope in the interior


def test_wave_equation_matches_the_reference_cpu_branch_in_float64():
    """Kamera.cs:213-218: v = v0 + n * 0.02*ctr*sin(40t + 100*sqrt(dx^2+dy^2)) — restated here in float64 numpy."""
    n = 4096
    rng = np.random.default_rng(3)
    base = rng.uniform(-1, 1, (n, 3)).astype(np.float32)
    nor = rng.uniform(-1, 1, (n, 3)).astype(np.float32)
    out = np.zeros((n, 3), np.float32)
    args = np.zeros(64, np.float32)
    args[:5] = [1.25, 0.3, 0.2, -0.4, n]
    po.wave_equation(base.reshape(-1), nor.reshape(-1), out.reshape(-1), args)
    dx, dy = base[:, 0].astype(np.float64) - np.float64(args[2]), base[:, 1].astype(np.float64) - np.float64(args[3])
    wave = 0.02 * 1.25 * np.sin(40.0 * np.float64(args[1]) + 100.0 * np.sqrt(dx * dx + dy * dy))
    assert np.allclose(out, base + nor * wave[:, None], rtol=0, atol=3e-6)
    args[4] = n // 2                                     # the kernel's own guard: work-items >= n leave the output alone
    out2 = np.full((n, 3), 7.0, np.float32)
    po.wave_equation(base.reshape(-1), nor.reshape(-1), out2.reshape(-1), args)
    assert np.array_equal(out2[: n // 2], out[: n // 2]) and (out2[n // 2:] == 7.0).all()


def test_splitmix_stream_is_deterministic():
    a, b = po.uniform(0xC0FFEE, 1000, -15, 15), po.uniform(0xC0FFEE, 1000, -15, 15)
    assert np.array_equal(a, b) and a.min() >= -15 and a.max() < 15 and abs(float(a.mean())) < 1.5
