"""Python binding with the reference's class names over libcekirdekler_host.so (which drives libKutuphaneCL.so).

Mirrors the C# surface of SURVEY.md Appendix F for the compute() path:

    Cekirdekler.AcceleratorType                         ClNumberCruncher.cs:32-49
    Cekirdekler.ClNumberCruncher                        ClNumberCruncher.cs:199-416
    Cekirdekler.ClArrays.ClArray<T> / ClParameterGroup  ClArray.cs:219-659, :717-1907
    Cekirdekler.ClArrays.Cl{Float,Double,Int,UInt,Long,Char,Byte}Array   CSpaceArrays.cs:582-1517
    Cekirdekler.Hardware.ClPlatforms / ClDevices        ClObjectApi.cs:204-1276

The arithmetic, scheduling and every transfer happen in the native libraries; this file only marshals.
numpy views give element access to the (pinned) native arrays, standing in for the C# indexers.
"""
from __future__ import annotations

import ctypes as C
import enum
from typing import Iterable, List, Optional, Sequence

import numpy as np

from ._native import host, shim

# CSpaceArrays.ARR_* (CSpaceArrays.cs:48-78); C# char is a UTF-16 code unit (2 bytes)
ARR_FLOAT, ARR_DOUBLE, ARR_LONG, ARR_INT, ARR_UINT, ARR_BYTE, ARR_CHAR = range(7)
_DTYPE_TO_ARR = {
    np.dtype(np.float32): ARR_FLOAT, np.dtype(np.float64): ARR_DOUBLE, np.dtype(np.int64): ARR_LONG,
    np.dtype(np.int32): ARR_INT, np.dtype(np.uint32): ARR_UINT, np.dtype(np.uint8): ARR_BYTE,
    np.dtype(np.uint16): ARR_CHAR,
}
_ARR_TO_DTYPE = {v: k for k, v in _DTYPE_TO_ARR.items()}

PIPELINE_EVENT = True    # Cores.cs:416-423
PIPELINE_DRIVER = False

_F_READ, _F_PARTIAL, _F_WRITE, _F_WRITEALL, _F_RO, _F_WO, _F_ZC, _F_ELEMS, _F_ALIGN = range(9)


class AcceleratorType(enum.IntFlag):
    CPU = 1
    GPU = 2
    ACC = 4


# ----------------------------------------------------------------------------- Hardware
class ClDevices:
    """Explicit device selection (ClObjectApi.cs:639-1276): indexer, Length, `+` concatenation, gpus()/cpus()/accelerators()."""

    def __init__(self, ordinals: Sequence[int] = (), streaming: bool = False):
        self.ordinals = list(ordinals)
        self.streamingEnabled = streaming

    @property
    def Length(self) -> int:
        return len(self.ordinals)

    def __len__(self):
        return len(self.ordinals)

    def __getitem__(self, i) -> "ClDevices":
        if isinstance(i, slice):
            return ClDevices(self.ordinals[i], self.streamingEnabled)
        return ClDevices([self.ordinals[i]], self.streamingEnabled)

    def __add__(self, other: "ClDevices") -> "ClDevices":  # ClObjectApi.cs:813-829
        return ClDevices(self.ordinals + other.ordinals, self.streamingEnabled or other.streamingEnabled)

    def gpus(self, devicePartitionEnabled: bool = False, streamingEnabled: bool = False, MAX_CPU: int = -1) -> "ClDevices":
        return ClDevices(self.ordinals, self.streamingEnabled or streamingEnabled)

    def cpus(self, *a, **k) -> "ClDevices":
        return ClDevices([])  # the CUDA platform has no CPU devices

    def accelerators(self, *a, **k) -> "ClDevices":
        return ClDevices([])

    def devicesNvidia(self) -> "ClDevices":
        return ClDevices(self.ordinals, self.streamingEnabled)

    def devicesAmd(self) -> "ClDevices":
        return ClDevices([])

    def devicesIntel(self) -> "ClDevices":
        return ClDevices([])

    def devicesWithMostComputeUnits(self) -> "ClDevices":
        h = host()
        return ClDevices(sorted(self.ordinals, key=lambda o: -h.ckh_gpu_compute_units(o)), self.streamingEnabled)

    def devicesWithHighestMemoryAvailable(self) -> "ClDevices":
        h = host()
        return ClDevices(sorted(self.ordinals, key=lambda o: -h.ckh_gpu_memory_size(o)), self.streamingEnabled)

    def devicesWithDedicatedMemory(self) -> "ClDevices":
        return ClDevices(self.ordinals, self.streamingEnabled)

    def devicesWithHostMemorySharing(self) -> "ClDevices":  # ClObjectApi.cs:1170: iGPUs/CPUs only; every B200 has its own HBM
        return ClDevices([], self.streamingEnabled)

    def devicesWithHighestIntrapolatedNbodyPerformance(self) -> "ClDevices":  # ClObjectApi.cs:1254-1257 throws as well
        raise NotImplementedError("not implemented in the reference either (ClObjectApi.cs:1256)")

    def devicesWithHighestLeastOscillatedNbodyPerformance(self) -> "ClDevices":  # ClObjectApi.cs:1267-1270
        raise NotImplementedError("not implemented in the reference either (ClObjectApi.cs:1269)")

    def devicesWithHighestDirectNbodyPerformance(self, n: int = 16 * 1024) -> "ClDevices":
        """Orders the devices by a direct n-body micro-benchmark, fastest first (ClObjectApi.cs:1222-1244 runs
        Tester.nBody(16K) on each device); here the 2-D reference kernel `nBody2D` at n bodies, timed on the device."""
        from . import kernels

        x = ClArray(n, np.float32)
        y = ClArray(n, np.float32)
        x.view[:] = (np.arange(n) * 0.37) % 30.0 - 15.0
        y.view[:] = (np.arange(n) * 0.73) % 30.0 - 15.0
        x.readOnly = True
        y.readOnly = True
        nn = ClArray(1, np.int32)
        nn[0] = n
        nn.readOnly = True
        timings = []
        for o in self.ordinals:
            cr = ClNumberCruncher(ClDevices([o]), kernels.source("nbody"))
            fx, fy = ClArray(n, np.float32), ClArray(n, np.float32)
            fx.writeOnly = True
            fy.writeOnly = True
            for _ in range(3):
                x.nextParam(y, fx, fy, nn).compute(cr, 1, "nBody2D", n, 256)
            timings.append((cr.benchmarks(1)[0] if cr.errorCode() == 0 else float("inf"), o))
            cr.dispose()
        return ClDevices([o for _, o in sorted(timings)], self.streamingEnabled)

    def names(self) -> List[str]:
        h = host()
        return [h.ckh_gpu_name(o).decode() for o in self.ordinals]

    def logInfo(self) -> str:  # ClObjectApi.cs:901-928
        h = host()
        lines = ["---------", "Selected devices:", "---------"]
        for k, o in enumerate(self.ordinals):
            lines.append(
                f"#{k}: {h.ckh_gpu_name(o).decode()}(NVIDIA Corporation) number of compute units: "
                f"{h.ckh_gpu_compute_units(o)} type:GPU memory: {h.ckh_gpu_memory_size(o) / 2**30:.2f}GB"
            )
        lines.append("---------")
        text = "\n".join(lines)
        print(text)
        return text


class ClPlatforms:
    """ClPlatforms.all() (ClObjectApi.cs:204): one synthetic platform, "NVIDIA CUDA"."""

    def __init__(self, n_gpus: int):
        self._n = n_gpus

    @staticmethod
    def all() -> "ClPlatforms":
        return ClPlatforms(host().ckh_gpu_count())

    @property
    def Length(self) -> int:
        return 1

    def __getitem__(self, i) -> "ClPlatforms":
        if i != 0:
            raise IndexError("only platform 0 (NVIDIA CUDA) exists")
        return self

    def platformVendorNames(self) -> List[str]:
        return ["NVIDIA Corporation"]

    def platformsNvidia(self) -> "ClPlatforms":
        return self

    def gpus(self, devicePartitionEnabled: bool = False, streamingEnabled: bool = False, MAX_CPU: int = -1) -> ClDevices:
        return ClDevices(range(self._n), streamingEnabled)

    def cpus(self, *a, **k) -> ClDevices:
        return ClDevices([])

    def accelerators(self, *a, **k) -> ClDevices:
        return ClDevices([])


# ----------------------------------------------------------------------------- arrays
class ClArray:
    """ClArray<T> (ClArray.cs:717-1907).  `ClArray(n, dtype)` owns a pinned native array (fastArr=True, the reference's
    default for `new ClArray<T>(n)`); `ClArray(ndarray)` wraps caller memory like the implicit conversion from `T[]`
    (ClArray.cs:1014-1046) — pageable unless `pin=True` page-locks it in place."""

    def __init__(self, source, dtype=np.float32, alignmentBytes: int = 4096, pin: bool = False):
        h = host()
        self._h = h
        self._keep = None
        self._registered = None
        if isinstance(source, (int, np.integer)):
            dt = np.dtype(dtype)
            self._handle = h.ckh_array_create(_DTYPE_TO_ARR[dt], int(source), alignmentBytes)
            if not self._handle:
                raise MemoryError("createArray failed")
            n = int(source)
        else:
            arr = np.ascontiguousarray(source)
            if arr is not source and not isinstance(source, np.ndarray):
                pass  # a copy of a Python sequence: results land in self.view
            dt = arr.dtype
            if dt not in _DTYPE_TO_ARR:
                raise TypeError(f"unsupported element type {dt}")
            self._keep = arr
            n = arr.size
            if pin and n:
                if shim().ckHostRegister(arr.ctypes.data, arr.nbytes) == 0:
                    self._registered = arr.ctypes.data
            self._handle = h.ckh_array_wrap(arr.ctypes.data, _DTYPE_TO_ARR[dt], n)
            if not self._handle:
                raise MemoryError("array wrap failed")
        self.dtype = dt
        self._n = n
        ptr = h.ckh_array_data(self._handle)
        if self._keep is not None:
            self.view = self._keep.reshape(-1)
        else:
            buf = (C.c_char * (n * dt.itemsize)).from_address(ptr) if n else (C.c_char * 0)()
            self.view = np.frombuffer(buf, dtype=dt, count=n)

    # -- lifetime
    def dispose(self):
        if getattr(self, "_handle", None):
            self._h.ckh_array_destroy(self._handle)
            self._handle = None
            if self._registered:
                shim().ckHostUnregister(self._registered)
                self._registered = None

    def __del__(self):
        try:
            self.dispose()
        except Exception:
            pass

    # -- element access (IList<T>, ClArray.cs:1896-1906)
    def __len__(self):
        return self._n

    @property
    def Length(self) -> int:
        return self._n

    @property
    def N(self) -> int:
        return self._n

    @N.setter
    def N(self, value: int):
        """Resize (ClArray.cs:749-800): a new array of the same kind, the first min(old, new) elements kept, the rest zero."""
        value = int(value)
        if not getattr(self, "_handle", None) or value == self._n:
            return
        self._replace_storage(value, self.fastArr)

    def _replace_storage(self, n: int, fast: bool):
        """New backing store of `n` elements (pinned native array if `fast`, a language array otherwise) holding the old
        values; flags, elements per work-item and alignment carry over.  Device buffers are keyed by the array object in
        the host layer, so the next compute() simply sees a new array."""
        old_view, keep_n = self.view, min(self._n, n)
        flags = [self._get(i) for i in range(9)]
        fresh = ClArray(n, self.dtype, alignmentBytes=max(flags[_F_ALIGN], 1)) if fast else ClArray(np.zeros(n, self.dtype))
        fresh.view[:keep_n] = old_view[:keep_n]
        fresh.view[keep_n:] = 0
        self.dispose()
        for name in ("_handle", "_keep", "_registered", "view", "_n"):
            setattr(self, name, getattr(fresh, name))
        fresh._handle, fresh._registered = None, None       # ownership moved
        # read/partial/write/writeAll first, then the exclusive switches (which clear the others), then the scalars
        for i in (_F_READ, _F_PARTIAL, _F_WRITE, _F_WRITEALL, _F_ZC, _F_ELEMS):
            self._set(i, flags[i])
        if flags[_F_RO]:
            self._set(_F_RO, 1)
        if flags[_F_WO]:
            self._set(_F_WO, 1)

    def __getitem__(self, i):
        return self.view[i]

    def __setitem__(self, i, v):
        self.view[i] = v

    def ToArray(self) -> np.ndarray:
        return self.view.copy()

    def CopyFrom(self, src, start: int = 0):
        src = np.asarray(src, dtype=self.dtype).reshape(-1)
        self.view[start:start + src.size] = src

    def CopyTo(self, dst: np.ndarray, start: int = 0):
        dst.reshape(-1)[start:start + self._n] = self.view

    @property
    def fastArr(self) -> bool:
        return bool(self._h.ckh_array_is_fast(self._handle))

    @fastArr.setter
    def fastArr(self, value: bool):
        """Switch between the C++ (pinned native) and the language-array representation, keeping the values
        (ClArray.cs:889-958; exercised by the reference's *ArrayOperations tests, Tester.cs:7470)."""
        if getattr(self, "_handle", None) and bool(value) != self.fastArr:
            self._replace_storage(self._n, bool(value))

    # -- flags (ClArray.cs:1742-1888)
    def _get(self, i):
        return self._h.ckh_array_get(self._handle, i)

    def _set(self, i, v):
        self._h.ckh_array_set(self._handle, i, int(v))

    read = property(lambda s: bool(s._get(_F_READ)), lambda s, v: s._set(_F_READ, v))
    partialRead = property(lambda s: bool(s._get(_F_PARTIAL)), lambda s, v: s._set(_F_PARTIAL, v))
    write = property(lambda s: bool(s._get(_F_WRITE)), lambda s, v: s._set(_F_WRITE, v))
    writeAll = property(lambda s: bool(s._get(_F_WRITEALL)), lambda s, v: s._set(_F_WRITEALL, v))
    readOnly = property(lambda s: bool(s._get(_F_RO)), lambda s, v: s._set(_F_RO, v))
    writeOnly = property(lambda s: bool(s._get(_F_WO)), lambda s, v: s._set(_F_WO, v))
    zeroCopy = property(lambda s: bool(s._get(_F_ZC)), lambda s, v: s._set(_F_ZC, v))
    numberOfElementsPerWorkItem = property(lambda s: s._get(_F_ELEMS), lambda s, v: s._set(_F_ELEMS, v))
    alignmentBytes = property(lambda s: s._get(_F_ALIGN), lambda s, v: s._set(_F_ALIGN, v))

    def flagString(self) -> str:
        return self._h.ckh_array_flag_string(self._handle).decode()

    # -- chaining and compute
    def nextParam(self, *arrays) -> "ClParameterGroup":
        g = ClParameterGroup()
        g._add(self, None)
        for a in arrays:
            g._add_any(a, raw_partial_default=False)  # ClArray.cs:1384-1386 (quirk Q11)
        return g

    @staticmethod
    def wrapArrayOfStructs(structs: np.ndarray) -> "ClArray":
        """ClArray<byte>.wrapArrayOfStructs (ClArray.cs:1058-1074): a byte view of an array of structs (for example an
        (n, 3) float32 array standing in for Vector3[]) with numberOfElementsPerWorkItem = the struct's size in bytes."""
        structs = np.ascontiguousarray(structs)
        item = structs.dtype.itemsize * int(np.prod(structs.shape[1:], dtype=np.int64)) if structs.ndim > 1 else structs.dtype.itemsize
        a = ClArray(structs.reshape(-1).view(np.uint8))
        a.numberOfElementsPerWorkItem = item
        a._structs = structs
        return a

    def compute(self, cruncher, computeId: int, kernelNamesString: str, globalRange: int, localRange: int = 256,
                ofsetGlobalRange: int = 0, pipeline: bool = False, pipelineType: bool = PIPELINE_EVENT, pipelineBlobs: int = 4,
                readWritesReady=None, elementsPerItemReady=None) -> int:
        g = ClParameterGroup()
        g._add(self, None)
        return g.compute(cruncher, computeId, kernelNamesString, globalRange, localRange, ofsetGlobalRange, pipeline,
                         pipelineType, pipelineBlobs, readWritesReady, elementsPerItemReady)


def _fast(dtype):
    def make(n: int, alignmentBytes: int = 4096) -> ClArray:
        return ClArray(int(n), dtype, alignmentBytes)
    return make


# concrete FastArr<T> types (CSpaceArrays.cs:582-1517): aligned native ("C++") arrays
ClFloatArray = _fast(np.float32)
ClDoubleArray = _fast(np.float64)
ClIntArray = _fast(np.int32)
ClUIntArray = _fast(np.uint32)
ClLongArray = _fast(np.int64)
ClCharArray = _fast(np.uint16)
ClByteArray = _fast(np.uint8)


class ClParameterGroup:
    """Ordered kernel-parameter list built by nextParam(); flags are snapshotted when an array is added, exactly as the
    reference copies them into its linked lists (ClArray.cs:219-500)."""

    def __init__(self):
        self._arrays: List[ClArray] = []
        self._flags: List[str] = []
        self._elems: List[int] = []

    def _add(self, a: ClArray, flags: Optional[str], elems: Optional[int] = None):
        self._arrays.append(a)
        self._flags.append(a.flagString() if flags is None else flags)
        self._elems.append(a.numberOfElementsPerWorkItem if elems is None else elems)

    def _add_any(self, a, raw_partial_default: bool):
        if isinstance(a, ClArray):
            self._add(a, None)
        else:
            # a raw array: wrapped with the defaults the reference gives C# arrays at this call site
            w = ClArray(a)
            self._add(w, (" partial " if raw_partial_default else "") + " read  write ", 1)

    def nextParam(self, *arrays) -> "ClParameterGroup":
        g = ClParameterGroup()
        g._arrays, g._flags, g._elems = list(self._arrays), list(self._flags), list(self._elems)
        for a in arrays:
            g._add_any(a, raw_partial_default=True)  # ClArray.cs:266-268 (quirk Q11)
        return g

    def compute(self, cruncher: "ClNumberCruncher", computeId: int, kernelNamesString: str, globalRange: int,
                localRange: int = 256, ofsetGlobalRange: int = 0, pipeline: bool = False, pipelineType: bool = PIPELINE_EVENT,
                pipelineBlobs: int = 4, readWritesReady=None, elementsPerItemReady=None) -> int:
        n = len(self._arrays)
        flags = readWritesReady if readWritesReady is not None else self._flags
        elems = elementsPerItemReady if elementsPerItemReady is not None else self._elems
        # the marshalled argument block is kept between calls: a ~50 us kernel is called thousands of times with the same
        # arrays and flags (SURVEY.md §7 hard part 6), and building four ctypes arrays costs several microseconds
        key = (tuple(flags), tuple(elems), kernelNamesString)
        handles = [a._handle for a in self._arrays]
        m = getattr(self, "_marshalled", None)
        if m is None or m[0] != key or m[1] != handles:
            m = (key, handles, (C.c_void_p * n)(*handles), (C.c_char_p * n)(*[f.encode() for f in flags]), (C.c_int32 * n)(*elems),
                 kernelNamesString.encode())
            self._marshalled = m
        return host().ckh_compute(cruncher._handle, m[2], n, computeId, m[5], globalRange, localRange,
                                  ofsetGlobalRange, int(pipeline), int(pipelineType), pipelineBlobs, m[3], m[4])


# ----------------------------------------------------------------------------- cruncher
class ClNumberCruncher:
    """ClNumberCruncher (ClNumberCruncher.cs:199-416).  `devices` is a ClDevices selection, or an AcceleratorType mask
    (only GPU devices exist on this platform; `numberOfGPUsToUse` limits how many)."""

    def __init__(self, devices, kernelString: str, noPipelining: bool = False, computeQueueConcurrency: int = 16,
                 numberOfGPUsToUse: int = -1, stream: Optional[bool] = None):
        h = host()
        self._h = h
        if isinstance(devices, ClDevices):
            ords = devices.ordinals
            streaming = devices.streamingEnabled if stream is None else stream
        else:
            mask = AcceleratorType(devices)
            ords = list(range(h.ckh_gpu_count())) if (mask & AcceleratorType.GPU) else []
            if numberOfGPUsToUse >= 0:
                ords = ords[:numberOfGPUsToUse]
            streaming = True if stream is None else stream  # the AcceleratorType ctor defaults stream=true (:201)
        arr = (C.c_int32 * max(1, len(ords)))(*ords)
        self._handle = h.ckh_cruncher_create(arr, len(ords), kernelString.encode(), int(noPipelining), computeQueueConcurrency,
                                             int(streaming))
        self._exchange_cb = None

    def dispose(self):
        if getattr(self, "_handle", None):
            self._h.ckh_cruncher_destroy(self._handle)
            self._handle = None

    def __del__(self):
        try:
            self.dispose()
        except Exception:
            pass

    def errorCode(self) -> int:
        return self._h.ckh_cruncher_error_code(self._handle)

    def errorMessage(self) -> str:
        return self._h.ckh_cruncher_error_message(self._handle).decode(errors="replace")

    @property
    def numberOfErrorsHappened(self) -> int:
        return self._h.ckh_cruncher_errors_happened(self._handle)

    @property
    def numberOfDevices(self) -> int:
        return self._h.ckh_cruncher_num_devices(self._handle)

    def deviceNames(self) -> List[str]:
        return [self._h.ckh_cruncher_device_name(self._handle, i).decode() for i in range(self.numberOfDevices)]

    @staticmethod
    def kernelNames(kernelString: str) -> List[str]:
        h = host()
        s = kernelString.encode()
        return [h.ckh_cruncher_kernel_name(s, i).decode() for i in range(h.ckh_cruncher_num_kernels(s))]

    def _prop(i):  # noqa: N805
        return property(lambda s: s._h.ckh_cruncher_get(s._handle, i), lambda s, v: s._h.ckh_cruncher_set(s._handle, i, int(v)))

    enqueueMode = _prop(0)
    enqueueModeAsyncEnable = _prop(1)
    noComputeMode = _prop(2)
    fineGrainedQueueControl = _prop(3)
    repeatCount = _prop(4)
    performanceFeed = _prop(5)
    smoothLoadBalancer = _prop(6)
    balancerMode = _prop(7)       # 0 = reference arithmetic, 1 = adaptive window (quirk Q14 fixed), 2 = cost-profile partition
    deviceClock = _prop(8)        # 1 = CUDA-event timing for the balancer, 0 = host stopwatch like the reference
    broadcastWholeReads = _prop(9)
    del _prop

    @property
    def repeatKernelName(self) -> str:
        return getattr(self, "_repeatKernelName", "")

    @repeatKernelName.setter
    def repeatKernelName(self, v: str):
        self._repeatKernelName = v
        self._h.ckh_cruncher_set_repeat_kernel_name(self._handle, v.encode())

    def countMarkersRemaining(self) -> int:
        return self._h.ckh_cruncher_markers_remaining(self._handle)

    def countMarkersReached(self) -> int:
        return self._h.ckh_cruncher_markers_reached(self._handle)

    def flushLastUsedCommandQueue(self):
        self._h.ckh_cruncher_flush_last_used_queue(self._handle)

    def sync(self):
        self._h.ckh_cruncher_sync(self._handle)

    def lastComputePerformanceReport(self, computeId: int = 0) -> str:
        return self._h.ckh_cruncher_performance_report(self._handle, computeId).decode()

    def peerTable(self, *targets: "ClArray") -> "ClArray":
        """(No reference counterpart.)  A small whole-`read` array for kernels that store into the OTHER devices' copies of
        the target arrays over NVLink (nBody3Dwp): element 0 = D, the number of devices of this cruncher; then, for each
        target in turn, the device address of its buffer on device 0..D-1.  The buffers are created here if compute() has
        not made them yet; the table stays valid until a target is resized or disposed.  One process driving all devices
        only: in one-process-per-GPU mode the other buffers live in other address spaces."""
        n = self.numberOfDevices
        table = ClArray(1 + max(n, 1) * max(len(targets), 1), np.int64)
        table.view[0] = n
        for t, target in enumerate(targets):
            for i in range(n):
                p = self._h.ckh_cruncher_buffer_pointer(self._handle, target._handle, i)
                if not p:
                    raise RuntimeError("peerTable: no device buffer for target %d on device %d" % (t, i))
                table.view[1 + t * n + i] = np.array([p], dtype=np.uint64).view(np.int64)[0]
        table.readOnly = True
        return table

    def performanceHistory(self, computeId: int) -> List[List[float]]:
        """Cores.performanceHistory (Cores.cs:1075-1098): the balancer's 10-deep window of normalised throughput shares,
        oldest row first, one column per device; all zero before the compute id has run."""
        buf = (C.c_double * (10 * 64))()
        d = self._h.ckh_cruncher_performance_history(self._handle, computeId, buf, len(buf))
        return [[buf[k * d + j] for j in range(d)] for k in range(10)] if d else []

    def _ivec(self, fn, computeId) -> List[int]:
        buf = (C.c_int32 * 64)()
        n = fn(self._handle, computeId, buf, 64)
        return list(buf[:n])

    def globalRanges(self, computeId: int) -> List[int]:
        return self._ivec(self._h.ckh_cruncher_ranges, computeId)

    def globalReferences(self, computeId: int) -> List[int]:
        return self._ivec(self._h.ckh_cruncher_references, computeId)

    def benchmarks(self, computeId: int) -> List[float]:
        buf = (C.c_double * 64)()
        n = self._h.ckh_cruncher_benchmarks(self._handle, computeId, buf, 64)
        return list(buf[:n])

    def normalizedComputePowersOfDevices(self) -> List[float]:
        buf = (C.c_double * 64)()
        n = self._h.ckh_cruncher_normalized_compute_powers(self._handle, buf, 64)
        return list(buf[:n])

    def normalizedGlobalRangesOfDevices(self, computeId: int) -> List[float]:
        buf = (C.c_double * 64)()
        n = self._h.ckh_cruncher_normalized_global_ranges(self._handle, computeId, buf, 64)
        return list(buf[:n])

    def resetBalancer(self, computeId: int):
        self._h.ckh_cruncher_reset_balancer(self._handle, computeId)

    def setSpmd(self, rank: int, world: int, exchange):
        """One process per GPU: `exchange(list_of_world_times) -> list_of_world_times` all-gathers the device times."""
        def cb(ptr, n, _user):
            vals = [ptr[i] for i in range(n)]
            out = exchange(vals)
            for i in range(n):
                ptr[i] = out[i]
        self._exchange_cb = self._h.EXCHANGE_FN(cb)
        self._h.ckh_cruncher_set_spmd(self._handle, rank, world, self._exchange_cb, None)
