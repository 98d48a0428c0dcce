"""ctypes loader for oracle/liboracle.so (the C++/OpenMP restatement in oracle.cpp).

TEST INFRASTRUCTURE ONLY — see oracle.cpp's header.  Parity status per function is stated there: the copy,
vectorAdd and 2-D n-body functions follow reference source lines; Mandelbrot, SAXPY, 3-D n-body, blur, Sobel and
histogram have no reference source or vectors ("parity unpinned") and are defined by this repo.
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

_DIR = Path(__file__).resolve().parent
_lib = None


def build(force: bool = False) -> Path:
    out = _DIR / "liboracle.so"
    srcs = sorted(_DIR.glob("*.cpp"))
    if force or not out.exists() or any(s.stat().st_mtime > out.stat().st_mtime for s in srcs):
        subprocess.run(
            ["g++", "-O3", "-march=x86-64-v3", "-ffp-contract=off", "-fno-fast-math", "-fopenmp", "-std=c++17", "-fPIC",
             "-Wall", "-shared", "-o", str(out)] + [str(s) for s in srcs], check=True)
    return out


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        L = C.CDLL(str(build()))
        vp, ll, i32 = C.c_void_p, C.c_longlong, C.c_int
        L.oracle_threads.restype = i32
        L.oracle_set_threads.argtypes = [i32]
        L.oracle_copy.argtypes = [vp, vp, ll, ll, i32]
        L.oracle_vector_add.argtypes = [vp, vp, vp, ll, ll]
        L.oracle_saxpy.argtypes = [vp, vp, vp, ll, ll]
        L.oracle_nbody2d.argtypes = [vp, vp, vp, vp, i32, ll, ll]
        L.oracle_mandelbrot.argtypes = [vp, vp, ll, ll]
        L.oracle_mandelbrot_sample.argtypes = [vp, ll, ll, ll, vp]
        L.oracle_mandelbrot_sample.restype = ll
        L.oracle_nbody3d_step.argtypes = [vp, vp, vp, vp, i32, ll, ll]
        L.oracle_blur5x5.argtypes = [vp, vp, vp, ll, ll]
        L.oracle_sobel.argtypes = [vp, vp, vp, ll, ll]
        L.oracle_histogram.argtypes = [vp, vp, ll, ll]
        L.oracle_wave_equation.argtypes = [vp, vp, vp, vp, ll, ll]
        _lib = L
    return _lib


def _p(a: np.ndarray):
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data


def threads() -> int:
    return lib().oracle_threads()


def use_all_cores() -> int:
    """Give the OpenMP runtime every host core this process may run on (launchers such as torchrun set OMP_NUM_THREADS=1)."""
    import os

    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    lib().oracle_set_threads(n)
    return threads()


def mandelbrot_view(W: int, H: int, maxIter: int, x0=-2.0, x1=1.0, y0=-1.5, y1=1.5) -> np.ndarray:
    """The `view` parameter array of the Mandelbrot kernel: {x0, y0, dx, dy, W, maxIter, 0, 0} in float32."""
    v = np.zeros(8, np.float32)
    v[0], v[1] = x0, y0
    v[2] = np.float32((x1 - x0) / W)
    v[3] = np.float32((y1 - y0) / H)
    v[4], v[5] = W, maxIter
    return v


def mandelbrot(view: np.ndarray, npix: int, offset: int = 0, count: int = -1, out: np.ndarray = None) -> np.ndarray:
    if out is None:
        out = np.zeros(npix, np.int32)
    lib().oracle_mandelbrot(_p(view), _p(out), offset, npix - offset if count < 0 else count)
    return out


def mandelbrot_sample(view: np.ndarray, offset: int, stride: int, count: int, out: np.ndarray = None) -> int:
    return lib().oracle_mandelbrot_sample(_p(view), offset, stride, count, _p(out) if out is not None else None)


def vector_add(a, b, c, offset=0, count=-1):
    lib().oracle_vector_add(_p(a), _p(b), _p(c), offset, a.size - offset if count < 0 else count)
    return c


def saxpy(alpha, x, y, offset=0, count=-1):
    lib().oracle_saxpy(_p(alpha), _p(x), _p(y), offset, x.size - offset if count < 0 else count)
    return y


def copy(src, dst, offset=0, count=-1):
    lib().oracle_copy(_p(src), _p(dst), offset, src.size - offset if count < 0 else count, src.dtype.itemsize)
    return dst


def nbody2d(x, y, fx, fy, offset=0, count=-1):
    lib().oracle_nbody2d(_p(x), _p(y), _p(fx), _p(fy), x.size, offset, x.size - offset if count < 0 else count)
    return fx, fy


def nbody3d_step(pos, vel, pos_out, params, offset=0, count=-1):
    n = pos.size // 4
    lib().oracle_nbody3d_step(_p(pos), _p(vel), _p(pos_out), _p(params), n, offset, n - offset if count < 0 else count)
    return pos_out


def blur5x5(img, out, dims, offset=0, count=-1):
    lib().oracle_blur5x5(_p(img), _p(out), _p(dims), offset, img.size - offset if count < 0 else count)
    return out


def sobel(img, out, dims, offset=0, count=-1):
    lib().oracle_sobel(_p(img), _p(out), _p(dims), offset, img.size - offset if count < 0 else count)
    return out


def histogram(img, hist, offset=0, count=-1):
    lib().oracle_histogram(_p(img), _p(hist), offset, img.size - offset if count < 0 else count)
    return hist


def wave_equation(xyz, xyzn, xyzo, arguments, offset=0, count=-1):
    n = xyz.size // 3
    lib().oracle_wave_equation(_p(xyz), _p(xyzn), _p(xyzo), _p(arguments), offset, n - offset if count < 0 else count)
    return xyzo


def splitmix64(seed: int, n: int) -> np.ndarray:
    """Deterministic uint64 stream (splitmix64) for seeded synthetic inputs (SURVEY.md §8c oracle plan)."""
    out = np.empty(n, np.uint64)
    x = np.uint64(seed)
    with np.errstate(over="ignore"):
        idx = np.arange(1, n + 1, dtype=np.uint64)
        z = x + idx * np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        out[:] = z ^ (z >> np.uint64(31))
    return out


def uniform(seed: int, n: int, lo: float, hi: float) -> np.ndarray:
    u = (splitmix64(seed, n) >> np.uint64(40)).astype(np.float64) / float(1 << 24)
    return (lo + (hi - lo) * u).astype(np.float32)
