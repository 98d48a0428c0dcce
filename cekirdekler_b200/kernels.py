"""Kernel source strings shipped with the package (what a user passes to ``ClNumberCruncher``).

The files under ``csrc/kernels/`` are hand-written CUDA for sm_100a in the ``__kernel void`` spelling the
reference's kernel-name regex recognises (ClNumberCruncher.cs:219-228).  At run time they go through the same
path as any user string: libKutuphaneCL.so rewrites the kernel signatures, prepends ``ck_prelude.cuh`` and
compiles with NVRTC for the device's architecture.  ``build()`` additionally compiles each of them offline with
``nvcc -gencode arch=compute_100a,code=sm_100a`` as a build check.
"""
from __future__ import annotations

import ctypes
from pathlib import Path

_DIR = Path(__file__).resolve().parent / "csrc" / "kernels"
_PRELUDE = Path(__file__).resolve().parent / "csrc" / "shim" / "ck_prelude.cuh"


def source(name: str) -> str:
    """Kernel source by file stem: 'mandelbrot', 'stream', 'nbody', 'image'.  A leading #line directive names the real
    file, so NVRTC's -lineinfo (and therefore ncu's source page) maps SASS back to csrc/kernels/<name>.cu."""
    path = _DIR / f"{name}.cu"
    return f'#line 1 "{path}"\n' + path.read_text()


def all_sources() -> dict:
    return {p.stem: p.read_text() for p in sorted(_DIR.glob("*.cu"))}


def nvrtc_options(src: str) -> list:
    """Options a source asks for through ``// ck-nvrtc: ...`` lines (same grammar as program.cpp)."""
    opts = []
    for line in src.splitlines():
        k = line.find("ck-nvrtc:")
        if k >= 0:
            opts += line[k + len("ck-nvrtc:"):].split()
    return opts


def rewritten_with_prelude(src: str) -> str:
    """The exact text NVRTC sees for ``src`` (prelude + rewritten kernels); used for the offline nvcc build."""
    from ._native import shim

    lib = shim()
    lib.ckRewriteKernelSource.restype = ctypes.c_void_p
    lib.ckRewriteKernelSource.argtypes = [ctypes.c_char_p]
    p = lib.ckRewriteKernelSource(src.encode())
    try:
        body = ctypes.string_at(p).decode()
    finally:
        ctypes.CDLL(None).free(ctypes.c_void_p(p))
    return _PRELUDE.read_text() + '\n#line 1 "user_kernels.cu"\n' + body
