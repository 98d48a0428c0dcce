"""Cekirdekler.Pipeline.Pool (ClPipeline.cs:3241-5078): tasks, task pools and the greedy device pool.

    ClTaskType       bit flags                                   :3247-3321
    ClTask           frozen compute() call, .type, .setCallBack  :3331-3520   (made by task(...) below = ClArray.task())
    ClTaskPool       feed / reset / remainingTaskGroupsOrTasks   :3607-3783
    ClDevicePoolType DEVICE_COMPUTE_AT_WILL, DEVICE_ROUND_ROBIN  :3802-3808
    ClDevicePool     addDevice / enqueueTaskPool / finish / dispose   :3872-4700
"""
from __future__ import annotations

import ctypes as C
import enum
from typing import List, Optional, Sequence

from ._native import host
from .api import PIPELINE_EVENT, ClArray, ClDevices, ClParameterGroup


class ClTaskType(enum.IntFlag):
    TASK_MESSAGE_DEFAULT = 0
    TASK_MESSAGE_DEVICE_SELECT_BEGIN = 1
    TASK_MESSAGE_DEVICE_SELECT_END = 2
    TASK_MESSAGE_GLOBAL_SYNCHRONIZATION_FIRST = 4
    TASK_MESSAGE_GLOBAL_SYNCHRONIZATION_LAST = 8
    TASK_MESSAGE_BROADCAST = 16
    TASK_MESSAGE_NO_COMPUTE = 32
    TASK_MESSAGE_SERIAL_MODE_BEGIN = 64
    TASK_MESSAGE_SERIAL_MODE_END = 128
    TASK_MESSAGE_LAST = 256


class ClDevicePoolType(enum.IntEnum):
    DEVICE_COMPUTE_AT_WILL = 0
    DEVICE_ROUND_ROBIN = 1


class ClTask:
    def __init__(self, data, computeId: int, kernelNamesString: str, globalRange: int, localRange: int = 256,
                 ofsetGlobalRange: int = 0, pipeline: bool = False, pipelineType: bool = PIPELINE_EVENT, pipelineBlobs: int = 4):
        self._h = host()
        if isinstance(data, ClArray):
            g = ClParameterGroup()
            g._add(data, None)
            data = g
        self._group = data                      # keeps the arrays alive
        n = len(data._arrays)
        arrs = (C.c_void_p * n)(*[a._handle for a in data._arrays])
        rw = (C.c_char_p * n)(*[f.encode() for f in data._flags])      # flags frozen now, as ClArray.task() does
        ep = (C.c_int32 * n)(*data._elems)
        self._handle = self._h.ckh_task_create(arrs, n, rw, ep, computeId, kernelNamesString.encode(), globalRange, localRange,
                                               ofsetGlobalRange, int(pipeline), int(pipelineType), pipelineBlobs)
        self._cb = None

    @property
    def type(self) -> ClTaskType:
        return ClTaskType(self._h.ckh_task_get_type(self._handle))

    @type.setter
    def type(self, v):
        self._h.ckh_task_set_type(self._handle, int(v))

    def setCallBack(self, fn):
        self._cb = self._h.CALLBACK_FN(lambda _user: fn())
        self._h.ckh_task_set_callback(self._handle, C.cast(self._cb, C.c_void_p), None)

    def compute(self, cruncher) -> int:
        return self._h.ckh_task_compute(self._handle, cruncher._handle)


def task(data, computeId: int, kernelNamesString: str, globalRange: int, localRange: int = 256, **kw) -> ClTask:
    """ClArray.task() / ClParameterGroup.task() (ClArray.cs:515-521, :1552-1583)."""
    return ClTask(data, computeId, kernelNamesString, globalRange, localRange, **kw)


class ClTaskPool:
    def __init__(self):
        self._h = host()
        self._handle = self._h.ckh_taskpool_create()
        self._tasks: List[ClTask] = []          # callbacks and arrays must outlive the native copies

    def feed(self, t: ClTask):
        self._tasks.append(t)
        self._h.ckh_taskpool_feed(self._handle, t._handle)

    def reset(self):
        self._h.ckh_taskpool_reset(self._handle)

    def remainingTaskGroupsOrTasks(self) -> int:
        return self._h.ckh_taskpool_remaining(self._handle)


class ClDevicePool:
    def __init__(self, poolType: ClDevicePoolType, kernelCodeToCompile: str, fineGrainedQueueControl: bool = False,
                 maxQueuesPerDevice: int = 3):
        self._h = host()
        self._handle = self._h.ckh_devicepool_create(int(poolType), kernelCodeToCompile.encode(), int(fineGrainedQueueControl),
                                                     maxQueuesPerDevice)
        self._pools = []

    def addDevice(self, devices: ClDevices):
        for o in devices.ordinals:
            self._h.ckh_devicepool_add_device(self._handle, o)

    def enqueueTaskPool(self, pool: ClTaskPool):
        self._pools.append(pool)
        self._h.ckh_devicepool_enqueue(self._handle, pool._handle)

    def finish(self) -> float:
        return self._h.ckh_devicepool_finish(self._handle)

    def errorCode(self) -> int:
        return self._h.ckh_devicepool_error_code(self._handle)

    def errorMessage(self) -> str:
        return self._h.ckh_devicepool_error_message(self._handle).decode(errors="replace")

    def tasksPerDevice(self) -> List[int]:
        buf = (C.c_int32 * 64)()
        n = self._h.ckh_devicepool_tasks_per_device(self._handle, buf, 64)
        return list(buf[:n])

    def dispose(self):
        if self._handle:
            self._h.ckh_devicepool_dispose(self._handle)
            self._h.ckh_devicepool_destroy(self._handle)
            self._handle = None
