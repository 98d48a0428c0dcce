"""The call sequence of the reference's C#, replayed against libKutuphaneCL.so through ctypes alone.

No CLR exists on either machine, so nothing here can load Cekirdekler.dll.  What the C# does at the native boundary is
known, though: Worker's constructor and its transfer/compute entry points (Worker.cs:36-65, :214-281, :821-1114,
:1308-1326, :1411-1567), the bulk waits of Cores (Cores.cs:39-49), the marker counters (ClNumberCruncher.cs:356-372,
ClCommandQueue.cs:31-47) and the user-event group (ClUserEvent.cs:30-46).  This file issues those calls in that order with
the handle discipline the C# has (events deleted right after they were handed to a wait list; deletes arriving late, twice,
in reverse and in random order; several threads at once as under Parallel.For), using only the 81 reference symbols, and
checks the results.  The last test asserts that every one of the 81 exports was called on the GPU by this file.
"""
import ctypes as C
import random
import threading
import time

import numpy as np
import pytest

from cekirdekler_b200._native import shim
from test_abi_symbols import REFERENCE_81

pytestmark = pytest.mark.gpu

CALLED = set()


class Recorder:
    """shim() with a note of which exports were called."""

    def __init__(self):
        self._L = shim()

    def __getattr__(self, name):
        fn = getattr(self._L, name)
        if name in REFERENCE_81:
            CALLED.add(name)
        return fn


L = None


@pytest.fixture(autouse=True)
def _lib(gpus):
    global L
    L = Recorder()
    yield


SRC = """
__kernel void scaleAdd(__global float* a, __global float* b)
{
    int i = get_global_id(0);
    b[i] = a[i] * 2.0f + 1.0f;
}
__kernel void addOne(__global float* a, __global float* b)
{
    int i = get_global_id(0);
    b[i] = b[i] + 1.0f;
}
__kernel void bump(__global float* a, __global float* b)
{
    if (get_global_id(0) == 0) a[0] = a[0] + 100.0f;
}
"""
N = 1 << 16
LOCAL = 256
ARR_FLOAT = 0        # CSpaceArrays.ARR_FLOAT (CSpaceArrays.cs:48-78)
CL_FLOAT = 0         # ClBuffer.SizeOf float (ClBuffer.cs:266-275)


def text(h):
    return L.readFromString(h).decode()


class RawWorker:
    """What `new Worker(device, source, kernelNames, ...)` does at the boundary (Worker.cs:214-281), then the arrays and
    buffers of one compute() call (Worker.cs:576-720, :990-1021)."""

    def __init__(self, index=0, partition=False, default_queue=False):
        self.plist = L.platformList()
        assert L.numberOfPlatforms(self.plist) >= 1
        self.plat = L.createPlatform(self.plist, 0)
        assert "CUDA" in text(L.getPlatformNameString(self.plat)) and "NVIDIA" in text(L.getPlatformVendorNameString(self.plat))
        L.getPlatformInfo(self.plat)                                        # imported by the C#, never called; must be harmless
        assert L.numberOfGpusInPlatform(self.plat) >= 1
        assert L.numberOfCpusInPlatform(self.plat) == 0 and L.numberOfAcceleratorsInPlatform(self.plat) == 0
        assert len({L.CODE_CPU(), L.CODE_GPU(), L.CODE_ACC()}) == 3
        self.dev = (L.createDeviceAsPartition(self.plat, L.CODE_GPU(), index, 4) if partition
                    else L.createDevice(self.plat, L.CODE_GPU(), index))
        assert self.dev
        name, vendor = L.createString(), L.createString()
        L.getDeviceName(self.dev, name)
        L.getDeviceVendorName(self.dev, vendor)
        self.name = text(name)
        assert self.name and "NVIDIA" in text(vendor)
        L.deleteString(name)
        L.jsonStringCallBack(vendor)                                        # "read done": releases a caller-made string as well
        assert L.deviceGDDR(self.dev) == 1 and L.deviceComputeUnits(self.dev) > 0 and L.deviceMemSize(self.dev) > (1 << 30)
        self.ctx = L.createContext(self.plat, self.dev)
        assert self.ctx
        self.q = L.createCommandQueue(self.ctx, self.dev, 0)                 # commandQueue
        self.q_read = L.createCommandQueue(self.ctx, self.dev, 0)            # commandQueueRead
        self.q_write = (L.createCommandQueue2(self.ctx, self.dev, 0, 1) if default_queue   # Worker.cs:224-227
                        else L.createCommandQueue(self.ctx, self.dev, 0))    # commandQueueWrite
        assert self.q and self.q_read and self.q_write
        src = L.writeToString(L.createString(), SRC.encode())
        self.program = L.createProgram(self.ctx, self.dev, src)
        assert L.getProgramErr(self.program) == 0, self.build_log()
        L.deleteString(src)
        self.kernels = {}
        for kname in ("scaleAdd", "addOne", "bump"):
            s = L.writeToString(L.createString(), kname.encode())
            k = L.createKernel(self.program, s)
            assert L.getKernelErr(k) == 0
            L.deleteString(s)
            self.kernels[kname] = k
        self.arr_a, self.arr_b = L.createArray(N, 4096, ARR_FLOAT), L.createArray(N, 4096, ARR_FLOAT)
        self.h_a, self.h_b = L.alignedArrHead(self.arr_a), L.alignedArrHead(self.arr_b)
        assert self.h_a % 4096 == 0 and self.h_b % 4096 == 0
        self.a = np.ctypeslib.as_array(C.cast(self.h_a, C.POINTER(C.c_float)), (N,))
        self.b = np.ctypeslib.as_array(C.cast(self.h_b, C.POINTER(C.c_float)), (N,))
        self.buf_a = L.createBuffer(self.ctx, N, CL_FLOAT, 0, self.h_a, 1, 0, 0)
        self.buf_b = L.createBuffer(self.ctx, N, CL_FLOAT, 0, self.h_b, 1, 0, 0)
        assert self.buf_a and self.buf_b
        for k in self.kernels.values():
            L.setKernelArgument(k, self.buf_a, 0)
            L.setKernelArgument(k, self.buf_b, 1)
        self.ranges = []

    def build_log(self):
        p = L.readProgramErrString(self.program)
        try:
            return C.string_at(p).decode() if p else ""
        finally:
            C.CDLL(None).free(C.c_void_p(p))                                # the CLR frees this string itself

    def rng(self, n):
        r = L.createNdRange(n)
        self.ranges.append(r)
        return r

    def handles(self):
        """(delete function, handle) of everything this worker made, children first."""
        out = [(L.deleteNdRange, r) for r in self.ranges]
        out += [(L.deleteKernel, k) for k in self.kernels.values()]
        out += [(L.deleteBuffer, self.buf_a), (L.deleteBuffer, self.buf_b), (L.deleteArray, self.arr_a), (L.deleteArray, self.arr_b),
                (L.deleteProgram, self.program), (L.deleteCommandQueue, self.q), (L.deleteCommandQueue, self.q_read),
                (L.deleteCommandQueue, self.q_write), (L.deleteContext, self.ctx), (L.deleteDevice, self.dev), (L.deletePlatform, self.plat),
                (L.deletePlatformList, self.plist)]
        return out

    def dispose(self, order="forward", twice=False, seed=0):
        L.wait3(self.q, self.q_read, self.q_write)
        self.a = self.b = None
        hs = self.handles()
        if order == "reverse":
            hs = hs[::-1]                                                   # parents first: children must keep them alive
        elif order == "random":
            random.Random(seed).shuffle(hs)
        for f, h in hs:
            f(h)
        if twice:
            for f, h in hs:
                f(h)                                                        # a finalizer running after dispose(): a no-op


def replay(w: RawWorker, rounds=3):
    """One worker's share of Cores.compute, plain and pipelined, `rounds` times."""
    off, glob, loc = w.rng(0), w.rng(N), w.rng(LOCAL)
    for r in range(rounds):
        base = float(r)
        w.a[:] = np.arange(N, dtype=np.float32) + base
        w.b[:] = -1.0
        # --- plain path: Worker.writeToBuffer / compute / readFromBuffer (Worker.cs:821-860, :1033-1060, :1343-1361)
        L.writeToBufferRanged(w.q, w.buf_a, 0, N, w.h_a)
        assert L.compute(w.q, w.kernels["scaleAdd"], off, glob, loc) == 0
        L.readFromBufferRanged(w.q, w.buf_b, 0, N, w.h_b)
        L.flush(w.q)
        L.finish(w.q)
        assert np.array_equal(w.b, w.a * 2 + 1)

        # --- event-driven pipeline, two blobs (Worker.cs:1411-1567, Cores.cs:1256-1366): upload on the read queue, kernel on
        #     the compute queue after it, download on the write queue after the kernel; the C# disposes each ClEvent as soon
        #     as it sits in the next stage's wait list (Cores.cs:1555)
        w.a[:] = np.arange(N, dtype=np.float32) * 0.5 + base
        w.b[:] = -1.0
        half = N // 2
        hglob = w.rng(half)
        for blob in range(2):
            ref = blob * half
            hoff = w.rng(ref)
            e_none, e_compute, e_write = L.createEventArr(0), L.createEventArr(0), L.createEventArr(1)
            e1, e2, e3 = L.createEvent(), L.createEvent(), L.createEvent()
            L.writeToBufferRangedEvent(w.q_read, w.buf_a, ref, half, w.h_a, e_none, e1)
            L.addToEventArr(e_compute, e1, 0)
            L.deleteEvent(e1)
            assert L.computeEvent(w.q, w.kernels["scaleAdd"], hoff, hglob, loc, e_compute, e2) == 0
            L.addToEventArr(e_write, e2, 1)
            L.deleteEvent(e2)
            L.readFromBufferRangedEvent(w.q_write, w.buf_b, ref, half, w.h_b, e_write, e3)
            L.deleteEvent(e3)
            for arr in (e_none, e_compute, e_write):
                L.deleteEventArr(arr)
        for q in (w.q_read, w.q, w.q_write):
            L.flush(q)
        L.wait3(w.q_read, w.q, w.q_write)
        assert np.array_equal(w.b, w.a * 2 + 1)

        # --- whole-array event forms + wait2 (ClBuffer.cs:37-256)
        w.a[:] = 7.0 + base
        w.b[:] = 0.0
        ea, eb = L.createEventArr(0), L.createEventArr(0)
        e1, e2 = L.createEvent(), L.createEvent()
        L.writeToBufferEvent(w.q_read, w.buf_a, w.h_a, ea, e1)
        L.addToEventArr(eb, e1, 0)
        assert L.computeEvent(w.q, w.kernels["scaleAdd"], off, glob, loc, eb, e2) == 0
        ec = L.createEventArr(0)
        L.addToEventArr(ec, e2, 0)
        L.readFromBufferEvent(w.q_write, w.buf_b, w.h_b, ec, None)
        L.wait2(w.q, w.q_write)
        assert (w.b == 15.0 + 2 * base).all()
        for e in (e1, e2):
            L.deleteEvent(e)
        for arr in (ea, eb, ec):
            L.deleteEventArr(arr)

        # --- repeats natively (Worker.cs:1062-1114): b += 1 ten times, with and without the sync kernel on a[0]
        L.writeToBuffer(w.q, w.buf_b, w.h_b)
        assert L.computeRepeated(w.q, w.kernels["addOne"], off, glob, loc, 10) == 0
        assert L.computeRepeatedWithSyncKernel(w.q, w.kernels["addOne"], off, glob, loc, 5, w.kernels["bump"], off) == 0
        L.readFromBuffer(w.q, w.buf_b, w.h_b)
        L.readFromBuffer(w.q, w.buf_a, w.h_a)
        qs = (C.c_void_p * 2)(w.q_read, w.q_write)
        L.waitN(qs, w.q, 2)                                                  # Worker.finishUsedComputeQueues (Worker.cs:364-420)
        assert (w.b == 30.0 + 2 * base).all()
        assert w.a[0] == 7.0 + base + 500.0 and (w.a[1:] == 7.0 + base).all()


def test_one_worker_replays_the_reference_call_sequence():
    w = RawWorker()
    replay(w)
    w.dispose("forward", twice=True)


def test_partition_device_and_default_queue_behave_like_plain_ones():
    """createDeviceAsPartition (CPU fission in the reference, ClDevice.cs:85-95) and createCommandQueue2's default device
    queue (Worker.cs:224-227) hand back working objects on a GPU."""
    w = RawWorker(partition=True, default_queue=True)
    replay(w, rounds=1)
    w.dispose("reverse")


def test_markers_count_and_reset():
    """fineGrainedQueueControl (ClNumberCruncher.cs:73-79, :356-372): one marker per enqueued compute; the counter reports how
    many have been reached and can be cleared."""
    w = RawWorker()
    off, glob, loc = w.rng(0), w.rng(N), w.rng(LOCAL)
    L.writeToBuffer(w.q, w.buf_a, w.h_a)
    for _ in range(6):
        assert L.compute(w.q, w.kernels["scaleAdd"], off, glob, loc) == 0
        L.addMarkerToCommandQueue(w.q)
    L.finish(w.q)
    assert L.getMarkerCounterOfCommandQueue(w.q) == 6
    L.resetMarkerCounterOfCommandQueue(w.q)
    assert L.getMarkerCounterOfCommandQueue(w.q) == 0
    L.addMarkerToCommandQueue(w.q)
    L.finish(w.q)
    assert L.getMarkerCounterOfCommandQueue(w.q) == 1
    w.dispose("random", seed=3)


def test_user_events_gate_a_queue_until_triggered():
    """ClUserEvent (ClUserEvent.cs:30-46): work enqueued behind addUserEvent does not start until the event fires, either
    directly (triggerUserEvent) or when a counter raised with incrementUserEvent has been decremented back to zero."""
    w = RawWorker()
    off, glob, loc = w.rng(0), w.rng(N), w.rng(LOCAL)
    w.a[:] = 3.0
    L.writeToBuffer(w.q, w.buf_a, w.h_a)
    L.finish(w.q)
    for mode in ("trigger", "counter"):
        w.b[:] = -5.0
        ue = L.createUserEvent(w.ctx)
        assert ue
        if mode == "counter":
            L.incrementUserEvent(ue)
            L.incrementUserEvent(ue)
        L.addUserEvent(w.q, ue)
        if b"not available" in shim().ckGetLastError():
            pytest.skip("stream memory-wait is not available in this driver")
        assert L.compute(w.q, w.kernels["scaleAdd"], off, glob, loc) == 0
        L.readFromBuffer(w.q, w.buf_b, w.h_b)
        L.flush(w.q)
        time.sleep(0.05)
        assert (w.b == -5.0).all()                                           # still gated
        if mode == "trigger":
            L.triggerUserEvent(ue)
        else:
            L.decrementUserEvent(ue, w.ctx)
            time.sleep(0.02)
            assert (w.b == -5.0).all()                                       # one of two decrements: still gated
            L.decrementUserEvent(ue, w.ctx)
        L.finish(w.q)
        assert (w.b == 7.0).all()
        L.deleteUserEvent(ue)
        L.deleteUserEvent(ue)
    w.dispose("reverse", twice=True)


def test_host_array_helpers():
    """sizeOf / copyMemory / setKernelArguments (CSpaceArrays.cs:108-147, ClKernel.cs:31-38)."""
    sizes = (C.c_int32 * 20)()
    L.sizeOf(sizes)
    assert list(sizes[:7]) == [4, 8, 8, 4, 4, 1, 2]
    w = RawWorker()
    w.a[:] = np.arange(N, dtype=np.float32)
    w.b[:] = 0
    L.copyMemory(w.h_b, w.h_a, N * 4)
    assert np.array_equal(w.b, w.a)
    L.setKernelArguments(w.kernels["scaleAdd"], w.buf_a, 0)                  # imported by the C#, never called
    off, glob, loc = w.rng(0), w.rng(N), w.rng(LOCAL)
    L.writeToBuffer(w.q, w.buf_a, w.h_a)
    assert L.compute(w.q, w.kernels["scaleAdd"], off, glob, loc) == 0
    L.readFromBuffer(w.q, w.buf_b, w.h_b)
    L.finish(w.q)
    assert np.array_equal(w.b, w.a * 2 + 1)
    w.dispose("random", seed=11, twice=True)


def test_build_errors_come_back_as_a_code_and_a_log():
    """ClProgram (ClProgram.cs:30-40, :64-101): a source that does not build gives a non-zero getProgramErr and a log the C#
    reads as a string (and frees itself); kernels of such a program report an error instead of launching."""
    w = RawWorker()
    assert "error" not in w.build_log()                  # a good build: the log is readable (and usually empty)
    src = L.writeToString(L.createString(), b"__kernel void broken(__global float* a) { a[0] = no_such_symbol; }")
    bad = L.createProgram(w.ctx, w.dev, src)
    L.deleteString(src)
    assert L.getProgramErr(bad) != 0
    p = L.readProgramErrString(bad)
    try:
        log = C.string_at(p).decode()
    finally:
        C.CDLL(None).free(C.c_void_p(p))
    assert "no_such_symbol" in log
    name = L.writeToString(L.createString(), b"broken")
    k = L.createKernel(bad, name)
    L.deleteString(name)
    assert L.getKernelErr(k) != 0
    off, glob, loc = w.rng(0), w.rng(N), w.rng(LOCAL)
    assert L.compute(w.q, k, off, glob, loc) != 0
    L.deleteKernel(k)
    L.deleteProgram(bad)
    w.dispose("forward")


def test_workers_on_several_threads_with_deletes_in_any_order(gpus):
    """Parallel.For bodies (Cores.cs:671-824): one thread per worker, all calling the library at once, each worker then torn
    down in a different order, some of them twice (dispose() followed by the finalizer)."""
    errors = []

    def body(t):
        try:
            w = RawWorker(index=t % gpus)
            replay(w, rounds=2)
            w.dispose(("forward", "reverse", "random")[t % 3], twice=(t % 2 == 0), seed=t)
        except BaseException as e:   # noqa: BLE001 - reported below with the thread's index
            errors.append((t, repr(e)))

    threads = [threading.Thread(target=body, args=(t,)) for t in range(6)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errors, errors


def test_zz_every_reference_export_was_called_on_the_gpu():
    missing = sorted(set(REFERENCE_81) - CALLED)
    assert not missing, missing
