"""Build recipes for the native libraries (nvcc for sm_100a, g++ for the host layer and the oracle).

Everything is built in-tree so the `.so` files travel to the GPU box with the snapshot:
  cekirdekler_b200/lib/libKutuphaneCL.so       the C-ABI CUDA shim (include/KutuphaneCL.h)
  cekirdekler_b200/lib/libcekirdekler_host.so  C++ host layer (include/cekirdekler_host.h), links the shim
  cekirdekler_b200/lib/kernels/*.cubin         offline nvcc builds of the shipped kernel sources (build check,
                                               `-Xptxas -v` / cuobjdump inspection; the run-time path is NVRTC)
  oracle/liboracle.so                          CPU restatement used only by tests/bench as the checker
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
PKG = ROOT / "cekirdekler_b200"
CSRC = PKG / "csrc"
LIB = PKG / "lib"
ORACLE = ROOT / "oracle"

NVCC = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
CUDA_HOME = Path(NVCC).resolve().parent.parent
GENCODE = ["-gencode", "arch=compute_100a,code=sm_100a"]


def _run(cmd, **kw):
    r = subprocess.run([str(c) for c in cmd], capture_output=True, text=True, **kw)
    if r.returncode != 0:
        sys.stderr.write("FAILED: " + " ".join(str(c) for c in cmd) + "\n" + r.stdout + r.stderr)
        raise RuntimeError("build step failed: " + " ".join(str(c) for c in cmd[:3]))
    return r.stdout + r.stderr


def _stale(target: Path, sources) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(Path(s).stat().st_mtime > t for s in sources)


def _embed_prelude() -> Path:
    src = CSRC / "shim" / "ck_prelude.cuh"
    out = CSRC / "shim" / "ck_prelude_embed.inc"
    text = 'R"CKPRELUDE(' + src.read_text() + ')CKPRELUDE"\n'
    if not out.exists() or out.read_text() != text:
        out.write_text(text)
    return out


def build_shim(force: bool = False) -> Path:
    LIB.mkdir(parents=True, exist_ok=True)
    inc = _embed_prelude()
    srcs = [CSRC / "shim" / "kutuphanecl.cu", CSRC / "shim" / "program.cpp"]
    extra = sorted((CSRC / "shim").glob("ext_*.cu"))
    deps = srcs + extra + [inc, CSRC / "shim" / "handles.h", ROOT / "include" / "KutuphaneCL.h", Path(__file__),
                           CSRC / "kernels" / "nbody.cu"]   # ext_noalloc.cu compiles the shipped n-body source statically
    out = LIB / "libKutuphaneCL.so"
    if force or _stale(out, deps):
        _run([NVCC, *GENCODE, "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC,-Wall", "-shared",
              "-cudart", "static", "-o", out, *srcs, *extra, "-lnvrtc", "-ldl", "-lpthread"])
    return out


def build_host(force: bool = False) -> Path:
    shim = build_shim()
    srcs = sorted((CSRC / "host").glob("*.cpp"))
    hdrs = sorted((CSRC / "host").glob("*.h")) + [ROOT / "include" / "cekirdekler_host.h",
                                                  ROOT / "include" / "KutuphaneCL.h"]
    out = LIB / "libcekirdekler_host.so"
    if force or _stale(out, srcs + hdrs + [shim, Path(__file__)]):
        _run(["g++", "-O2", "-g", "-std=c++17", "-fPIC", "-Wall", "-shared", "-pthread", "-I", ROOT / "include",
              "-o", out, *srcs, "-L", LIB, "-lKutuphaneCL", "-Wl,-rpath,$ORIGIN"])
    return out


def kernel_sources():
    return sorted((CSRC / "kernels").glob("*.cu"))


def build_kernel_cubins(force: bool = False, verbose: bool = False):
    """Offline nvcc build of every shipped kernel source, through the same prelude+rewrite the run-time uses."""
    from . import kernels as K  # lazy: needs the shim for the rewrite

    outdir = LIB / "kernels"
    outdir.mkdir(parents=True, exist_ok=True)
    logs = {}
    for src in kernel_sources():
        cubin = outdir / (src.stem + ".cubin")
        gen = outdir / (src.stem + ".gen.cu")
        deps = [src, CSRC / "shim" / "ck_prelude.cuh", CSRC / "shim" / "program.cpp", Path(__file__)]
        if not (force or _stale(cubin, deps)):
            continue
        text = K.rewritten_with_prelude(src.read_text())
        gen.write_text(text)
        opts = K.nvrtc_options(src.read_text())
        nvcc_opts = []
        for o in opts:
            if o.startswith("--fmad=") or o.startswith("-fmad="):
                nvcc_opts.append("--fmad=" + o.split("=")[1])
            elif o in ("--use_fast_math", "-use_fast_math"):
                nvcc_opts.append("--use_fast_math")
        logs[src.stem] = _run([NVCC, *GENCODE, "-lineinfo", "-O3", "-std=c++17", "-cubin", "-Xptxas", "-v",
                               *nvcc_opts, "-o", cubin, gen])
        if verbose:
            print(logs[src.stem])
    return logs


def build_oracle(force: bool = False) -> Path:
    srcs = sorted(ORACLE.glob("*.cpp"))
    hdrs = sorted(ORACLE.glob("*.h"))
    out = ORACLE / "liboracle.so"
    if force or _stale(out, srcs + hdrs + [Path(__file__)]):
        # -ffp-contract=off: no silent FMA fusion — every fused operation in the oracle is an explicit fmaf()
        _run(["g++", "-O3", "-march=x86-64-v3", "-ffp-contract=off", "-fno-fast-math", "-fopenmp", "-std=c++17",
              "-fPIC", "-Wall", "-shared", "-o", out, *srcs])
    return out


def build_all(force: bool = False, verbose: bool = False):
    build_shim(force)
    build_host(force)
    build_oracle(force)
    build_kernel_cubins(force, verbose)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose=True)
