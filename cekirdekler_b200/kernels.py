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


# ---- reference kernel strings, verbatim OpenCL C (used by the parity tests) ----

def reference_copy_kernel(ctype: str) -> str:
    """The `test` copy kernel of the reference's Tester matrix for one element type (Tester.cs:35-38, :87-91 ...)."""
    return (
        "__kernel void test(__global " + ctype + " * data,__global " + ctype + " * data2)\n"
        "{\n"
        "    int id=get_global_id(0);\n"
        "    data2[id]=data[id];\n"
        "}\n"
    )


# C# char (2 bytes) is copied as two uchars per work-item (Tester.cs:117-121)
REFERENCE_COPY_KERNEL_CHAR = """
                __kernel void test(__global uchar * data,__global uchar * data2)
                { 
                    data2[get_global_id(0)*2]=data[get_global_id(0)*2];  
                    data2[get_global_id(0)*2+1]=data[get_global_id(0)*2+1];  
                }"""

REFERENCE_VECTOR_ADD = """
                __kernel void vectorAdd(__global float*a,__global float*b,__global float*c)
                {
                    int i=get_global_id(0);
                    c[i]=a[i]+b[i];
                }"""


def reference_nbody_kernel(n: int) -> str:
    """The reference's nBody kernel text with n spliced in, as Tester.cs:7727-7742 builds it."""
    return """
                __kernel void nBody(__global float * x,__global float * y,__global float * fx,__global float * fy)
                {
                    int i=get_global_id(0);    
                    float fx0 = 0;
                    float fy0 = 0;
                    for (int j = 0; j < """ + str(n) + """; j++)
                    {
                        float dx = x[i] - x[j];
                        float dy = y[i] - y[j];
                        float r = (float)sqrt(dx*dx+dy*dy+0.0001f);
                        fx0 += (dx * (1.0f / (r * r * r)));
                        fy0 += (dx * (1.0f / (r * r * r)));
                    }
                    fx[i] = fx0;
                    fy[i] = fy0;
                }
                """


# The Unity sample's kernel, verbatim (/root/reference/Kamera.cs:234-245): Vector3[] arrays arrive as byte arrays wrapped with
# ClArray<byte>.wrapArrayOfStructs (numberOfElementsPerWorkItem = sizeof(Vector3) = 12) and are read as floats by the kernel.
REFERENCE_WAVE_EQUATION = """
                        __kernel void waveEquation( __global float *xyz,__global float *xyzn,__global float *xyzo,__global float * arguments)
                        {
                            int threadId=get_global_id(0);
                            if(threadId<arguments[4])
                            {
                                float dx=xyz[threadId*3]-arguments[2];float dy=xyz[threadId*3+1]-arguments[3];float t=arguments[1];
                                float ctr=arguments[0];float wave=0.02f*ctr*sin(40.0f*t+100.0f*sqrt(dx*dx+dy*dy));
                                xyzo[threadId*3]=xyz[threadId*3]+xyzn[threadId*3]*wave; // wave equation for all surface vertices
                                xyzo[threadId*3+1]=xyz[threadId*3+1]+xyzn[threadId*3+1]*wave; // wave equation for all surface vertices
                                xyzo[threadId*3+2]=xyz[threadId*3+2]+xyzn[threadId*3+2]*wave; // wave equation for all surface vertices
                            }
                        }
                    """
