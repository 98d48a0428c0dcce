"""cekirdekler_b200 — B200-native implementation of Cekirdekler's ClNumberCruncher / Cores.compute() hot path.

Layers:
  lib/libKutuphaneCL.so        the reference's 81-symbol native boundary on CUDA (include/KutuphaneCL.h)
  lib/libcekirdekler_host.so   C++ host layer mirroring the C# classes (include/cekirdekler_host.h)
  api.py                       Python binding with the reference's class names
  csrc/kernels/*.cu            hand-written sm_100a kernels for the benchmark workloads
"""
from .api import (  # noqa: F401
    ARR_BYTE, ARR_CHAR, ARR_DOUBLE, ARR_FLOAT, ARR_INT, ARR_LONG, ARR_UINT, PIPELINE_DRIVER, PIPELINE_EVENT,
    AcceleratorType, ClArray, ClByteArray, ClCharArray, ClDevices, ClDoubleArray, ClFloatArray, ClIntArray,
    ClLongArray, ClNumberCruncher, ClParameterGroup, ClPlatforms, ClUIntArray,
)
from . import kernels  # noqa: F401
