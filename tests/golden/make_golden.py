"""Generates the committed golden fixtures under tests/golden/ from the CPU oracle.

Run from the repo root:  python tests/golden/make_golden.py
The reference holds no golden vectors for this path (no seeds, no golden files) and cannot be run here
(no CLR, no OpenCL), so these fixtures pin the ORACLE (and through it the CUDA path) against drift; they do not pin
the oracle to the reference.  Where the reference does fix a value — the copy matrix, vectorAdd c[700] — the tests
check that value directly instead.
"""
import hashlib
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from oracle import balancer as ob  # noqa: E402
from oracle import pyoracle as po  # noqa: E402

OUT = Path(__file__).resolve().parent


def balancer_cases():
    rng = np.random.default_rng(0xC0FFEE)
    cases = []
    for devices, total, local, smooth, pipeline, blobs in [
        (2, 1024, 64, True, False, 4), (2, 1024, 64, False, False, 4), (3, 8192, 64, True, False, 4),
        (8, 1 << 20, 256, True, False, 4), (8, 1 << 20, 256, False, True, 16), (4, 1 << 22, 256, True, True, 16),
    ]:
        bal = ob.ComputeIdBalancer(devices)
        speed = rng.uniform(0.5, 2.0, devices)  # relative device speeds
        calls = []
        bench = [0.0] * devices
        for _ in range(25):
            r, _ = bal.step(bench, smooth, total, pipeline, blobs, local)
            calls.append({"bench": [float(b) for b in bench], "ranges": r})
            bench = [float(round(r[i] / speed[i] / total * 100.0 * (1.0 + 0.02 * rng.standard_normal()), 6)) for i in range(devices)]
            bench = [max(b, 0.0) for b in bench]
        cases.append({"devices": devices, "total": total, "local": local, "smooth": smooth, "pipeline": pipeline,
                      "blobs": blobs, "calls": calls})
    (OUT / "balancer_cases.json").write_text(json.dumps(cases))


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def kernel_vectors():
    g = {}
    # Mandelbrot: C1 (1024^2, 256 iterations) full hash + a 64x64 crop kept verbatim; C3 strided sample
    v1 = po.mandelbrot_view(1024, 1024, 256)
    m1 = po.mandelbrot(v1, 1024 * 1024)
    g["mandelbrot_c1"] = {"view": v1.tolist(), "sha256": sha(m1), "sum": int(m1.sum()),
                          "crop_rows_480_544_cols_200_264": m1.reshape(1024, 1024)[480:544, 200:264].tolist()}
    v3 = po.mandelbrot_view(16384, 16384, 4096)
    s3 = np.zeros(65536, np.int32)
    total = po.mandelbrot_sample(v3, 12345, 4093, 65536, s3)
    g["mandelbrot_c3_sample"] = {"view": v3.tolist(), "offset": 12345, "stride": 4093, "count": 65536,
                                 "sha256": sha(s3), "sum": int(total)}
    # n-body 2-D (reference-faithful), n = 1024, seeded
    n = 1024
    x, y = po.uniform(0xC0FFEE, n, -15, 15), po.uniform(0xC0FFEE + 1, n, -15, 15)
    fx, fy = np.zeros(n, np.float32), np.zeros(n, np.float32)
    po.nbody2d(x, y, fx, fy)
    g["nbody2d_1024"] = {"seed_x": 0xC0FFEE, "seed_y": 0xC0FFEE + 1, "sha256_fx": sha(fx), "fx_head": fx[:8].tolist()}
    # n-body 3-D step, n = 2048
    n = 2048
    pos = np.zeros((n, 4), np.float32)
    for c in range(3):
        pos[:, c] = po.uniform(0xC0FFEE + 10 + c, n, -15, 15)
    pos[:, 3] = 1.0
    vel = np.zeros((n, 4), np.float32)
    out = np.zeros((n, 4), np.float32)
    po.nbody3d_step(pos.reshape(-1), vel.reshape(-1), out.reshape(-1), np.array([1e-3, 1e-4, n, 0], np.float32))
    g["nbody3d_2048"] = {"seed": 0xC0FFEE + 10, "pos_out_head": out[:4].tolist(), "vel_head": vel[:4].tolist()}
    # image chain 256 x 128
    W, H = 256, 128
    img = po.uniform(0xF00D, W * H, 0, 1)
    dims = np.array([W, H], np.int32)
    b, s = np.zeros_like(img), np.zeros_like(img)
    hist = np.zeros(256, np.uint32)
    po.blur5x5(img, b, dims)
    po.sobel(b, s, dims)
    po.histogram(s, hist)
    g["image_256x128"] = {"seed": 0xF00D, "sha256_blur": sha(b), "sha256_sobel": sha(s), "hist": hist.tolist()}
    (OUT / "kernel_vectors.json").write_text(json.dumps(g))


if __name__ == "__main__":
    balancer_cases()
    kernel_vectors()
    print("wrote", sorted(p.name for p in OUT.glob("*.json")))
