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


def strip_launch(W: int, H: int, rows_per_band: int = 30, local: int = 256):
    """(global range, local range) for the strip kernels of image.cu (blurStrips, sobelStrips, blurSobelHistogram): one warp
    per strip of 120 columns and band of `rows_per_band` rows, `local // 32` strips of one band per work-group.
    30 rows per band gives 288 work-groups for a 3840 x 2160 frame: two per SM on a B200's 148 SMs."""
    strips = (W + 119) // 120
    wpc = local // 32
    groups = ((strips + wpc - 1) // wpc) * ((H + rows_per_band - 1) // rows_per_band)
    return groups * local, local


def image_pipeline_source() -> str:
    """image.cu plus the three stage kernels of the blur -> Sobel -> histogram pipeline (BASELINE.json configs[4]) with the
    parameter order ClPipelineStage gives its kernels: inputs, hidden buffers, outputs (ClPipeline.cs:281-370).  Each stage is
    the strip kernel of its step; `dims` = { W, H, rows per band } travels down the pipeline beside the frame (the API has
    no scalar arguments), and a stage that has not received a frame yet (dims still zero) does nothing."""
    return source("image") + """
// ck-launch-bounds: blurStage 256 2
__kernel void blurStage(const float* __restrict__ in, const int* __restrict__ dims, float* __restrict__ out, int* __restrict__ dimsOut)
{
    if (get_global_id(0) < 3) dimsOut[get_global_id(0)] = dims[get_global_id(0)];
    ck_strips<true, false, false>(in, out, nullptr, nullptr, dims, (unsigned)ck_global_group_id(), 0u);
}
// ck-launch-bounds: sobelStage 256 2
__kernel void sobelStage(const float* __restrict__ in, const int* __restrict__ dims, float* __restrict__ out, int* __restrict__ dimsOut)
{
    if (get_global_id(0) < 3) dimsOut[get_global_id(0)] = dims[get_global_id(0)];
    ck_strips<false, true, false>(in, out, nullptr, nullptr, dims, (unsigned)ck_global_group_id(), 0u);
}
// one launch: counts, publishes into `hist` (overwriting it) and re-arms `acc`, a hidden buffer of the stage
// ck-launch-bounds: histStage 256
__kernel void histStage(const float* __restrict__ in, const int* __restrict__ dims, unsigned int* __restrict__ acc, unsigned int* __restrict__ hist)
{
    histogramStripsBody(in, hist, dims, acc);
}
// the whole chain as ONE stage on one device: the two intermediate frames never exist
// ck-launch-bounds: fusedStage 256 2
__kernel void fusedStage(const float* __restrict__ in, const int* __restrict__ dims, unsigned int* __restrict__ acc, unsigned int* __restrict__ hist)
{
    ck_strips<true, true, true>(in, nullptr, acc, hist, dims, (unsigned)get_group_id(0), (unsigned)get_num_groups(0));
}
"""
