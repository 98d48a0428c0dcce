"""Cekirdekler.Pipeline.SingleGPUPipeline (ClPipeline.cs:2357-3240): N kernel stages overlapped on ONE device.

    DevicePipeline(ClDevices, kernelCodes, computeQueueConcurrency=16)   :2369
    DevicePipelineStage(kernelNames, globalRange, localRange)            :3098   .bindArray(...)  .stopHostDeviceTransmission
    DevicePipelineArray(type, array)                                     :3186
    DevicePipelineArrayType { INPUT, OUTPUT, INTERNAL, TRANSITION }      :3178-3184

Every stage is enqueued on its own stream; a TRANSITION array is double-buffered in HBM so that, in parallel mode, stage k
consumes what stage k-1 produced during the previous feed() and all stages of one feed() overlap.  `latency` says how many
feeds lie between handing an input in and reading its result (0 in serial mode, stages-1 in parallel mode).
"""
from __future__ import annotations

import enum

import numpy as np

from ._native import host
from .api import _DTYPE_TO_ARR, ClArray, ClDevices


class DevicePipelineArrayType(enum.IntEnum):
    INPUT = 0
    OUTPUT = 1
    INTERNAL = 2
    TRANSITION = 3


class DevicePipelineArray:
    """INPUT/OUTPUT: wraps a ClArray (or numpy array) the caller reads/writes.  INTERNAL/TRANSITION: give (dtype, n[, elems])."""

    def __init__(self, type: DevicePipelineArrayType, array):
        self._h = host()
        self.type = DevicePipelineArrayType(type)
        self._keep = None
        if self.type in (DevicePipelineArrayType.INPUT, DevicePipelineArrayType.OUTPUT):
            if not isinstance(array, ClArray):
                array = ClArray(array)
            self._keep = array
            self.array = array
            self._handle = self._h.ckh_dparray_wrap(int(self.type), array._handle)
        else:
            if isinstance(array, ClArray):
                dt, n, e = array.dtype, array.Length, array.numberOfElementsPerWorkItem
            elif isinstance(array, np.ndarray):
                dt, n, e = array.dtype, array.size, 1
            else:
                dt, n = np.dtype(array[0]), int(array[1])
                e = int(array[2]) if len(array) > 2 else 1
            self.array = None
            self._handle = self._h.ckh_dparray_create(int(self.type), _DTYPE_TO_ARR[np.dtype(dt)], n, e)


class DevicePipelineStage:
    def __init__(self, kernelNames: str, globalRange: int, localRange: int):
        self._h = host()
        self._handle = self._h.ckh_dpstage_create(kernelNames.encode(), globalRange, localRange)
        self._arrays = []

    def bindArray(self, array: DevicePipelineArray):
        self._arrays.append(array)
        self._h.ckh_dpstage_bind_array(self._handle, array._handle)

    @property
    def stopHostDeviceTransmission(self) -> bool:
        return getattr(self, "_stop", False)

    @stopHostDeviceTransmission.setter
    def stopHostDeviceTransmission(self, v: bool):
        self._stop = bool(v)
        self._h.ckh_dpstage_stop_transmission(self._handle, int(v))


class DevicePipeline:
    def __init__(self, selectedDevice: ClDevices, kernelCodesC99: str, computeQueueConcurrency: int = 16):
        self._h = host()
        self._handle = self._h.ckh_dpipeline_create(selectedDevice.ordinals[0], kernelCodesC99.encode(), computeQueueConcurrency)
        self._stages = []

    def addStage(self, stage: DevicePipelineStage):
        self._stages.append(stage)
        self._h.ckh_dpipeline_add_stage(self._handle, stage._handle)

    def feed(self):
        self._h.ckh_dpipeline_feed(self._handle)

    def feedAsync(self, fn):
        """Runs `fn()` on the host while the device works on this feed (ClPipeline.cs:2584-2601)."""
        self._h.ckh_dpipeline_feed_async_begin(self._handle)
        fn()
        self._h.ckh_dpipeline_feed_async_end(self._handle)

    def feedAsyncBegin(self):
        self._h.ckh_dpipeline_feed_async_begin(self._handle)

    def feedAsyncEnd(self):
        self._h.ckh_dpipeline_feed_async_end(self._handle)

    def enableSerialMode(self):
        self._h.ckh_dpipeline_set_serial(self._handle, 1)

    def enableParallelMode(self):
        self._h.ckh_dpipeline_set_serial(self._handle, 0)

    def errorCode(self) -> int:
        return self._h.ckh_dpipeline_error_code(self._handle)

    def errorMessage(self) -> str:
        return self._h.ckh_dpipeline_error_message(self._handle).decode(errors="replace")

    @property
    def latency(self) -> int:
        return self._h.ckh_dpipeline_latency(self._handle)
