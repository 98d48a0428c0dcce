"""Pipeline namespace of the reference (ClPipeline.cs): device-to-device stage pipeline.

    Cekirdekler.Pipeline.ClPipelineStage   ClPipeline.cs:146-1873
    Cekirdekler.Pipeline.ClPipeline        ClPipeline.cs:41-139

Stage buffers live in HBM; a hand-off between stages is one NVLink peer copy (`hostForwarding=False`, default) instead of
the reference's download + host memcpy + upload (`hostForwarding=True` keeps that path for comparison).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np

from ._native import host
from .api import _DTYPE_TO_ARR, ClDevices


class ClPipelineStage:
    def __init__(self, debugLog: bool = False):
        self._h = host()
        self._handle = self._h.ckh_stage_create(int(debugLog))
        self._pipelines = []

    def addDevices(self, devices: ClDevices):
        o = devices.ordinals
        self._h.ckh_stage_add_devices(self._handle, (C.c_int32 * max(1, len(o)))(*o), len(o))

    def addKernels(self, kernels: str, kernelNames: str, globalRanges: Sequence[int], localRanges: Sequence[int]):
        n = len(globalRanges)
        self._h.ckh_stage_add_kernels(self._handle, kernels.encode(), kernelNames.encode(), (C.c_int32 * n)(*globalRanges),
                                      (C.c_int32 * n)(*localRanges), n)

    def initializerKernel(self, kernelNames: str, globalRanges: Sequence[int], localRanges: Sequence[int]):
        n = len(globalRanges)
        self._h.ckh_stage_initializer_kernel(self._handle, kernelNames.encode(), (C.c_int32 * n)(*globalRanges),
                                             (C.c_int32 * n)(*localRanges), n)

    def _add(self, role, specs):
        # a spec is a numpy array (its dtype/size describe the buffer), or (dtype, n[, elementsPerWorkItem])
        for s in specs:
            if isinstance(s, np.ndarray):
                dt, n, e = s.dtype, s.size, 1
            else:
                dt, n = np.dtype(s[0]), int(s[1])
                e = int(s[2]) if len(s) > 2 else 1
            self._h.ckh_stage_add_buffer(self._handle, role, _DTYPE_TO_ARR[np.dtype(dt)], n, e)

    def addInputBuffers(self, *specs):
        self._add(0, specs)

    def addOutputBuffers(self, *specs):
        self._add(1, specs)

    def addHiddenBuffers(self, *specs):
        self._add(2, specs)

    def prependToStage(self, stage: "ClPipelineStage"):
        self._h.ckh_stage_prepend_to(self._handle, stage._handle)

    def appendToStage(self, stage: "ClPipelineStage"):
        self._h.ckh_stage_append_to(self._handle, stage._handle)

    @property
    def enqueueMode(self) -> bool:
        return getattr(self, "_enq", False)

    @enqueueMode.setter
    def enqueueMode(self, v: bool):
        self._enq = bool(v)
        self._h.ckh_stage_set_enqueue_mode(self._handle, int(v))

    def makePipeline(self, hostForwarding: bool = False) -> "ClPipeline":
        p = ClPipeline(self._h.ckh_stage_make_pipeline(self._handle, int(hostForwarding)), self)
        return p


class ClPipeline:
    def __init__(self, handle, keep):
        self._h = host()
        self._handle = handle
        self._keep = keep  # stages must outlive the pipeline

    def errorCode(self) -> int:
        return self._h.ckh_pipeline_error_code(self._handle)

    def errorMessage(self) -> str:
        return self._h.ckh_pipeline_error_message(self._handle).decode(errors="replace")

    @property
    def peerToPeer(self) -> bool:
        return self._h.ckh_pipeline_forwarding(self._handle) == 0

    @property
    def counter(self) -> int:
        return self._h.ckh_pipeline_counter(self._handle)

    def pushData(self, data: Optional[Sequence[np.ndarray]] = None, popResultsHere: Optional[Sequence[np.ndarray]] = None) -> bool:
        def ptrs(arrs):
            if not arrs:
                return None, 0
            for a in arrs:
                assert a.flags["C_CONTIGUOUS"]
            return (C.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs]), len(arrs)
        d, nd = ptrs(data)
        r, nr = ptrs(popResultsHere)
        return bool(self._h.ckh_pipeline_push(self._handle, d, nd, r, nr))

    def dispose(self):
        if self._handle:
            self._h.ckh_pipeline_destroy(self._handle)
            self._handle = None
