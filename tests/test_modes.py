"""The execution modes of ClNumberCruncher that the pipelines and pools are built from (Cores.cs:60-127,
ClNumberCruncher.cs:66-187): enqueueMode, enqueueModeAsyncEnable, fineGrainedQueueControl markers, noComputeMode,
repeatCount with a sync kernel, zero-copy ("stream") arrays, plus the validation rules of ClArray.compute (ClArray.cs:1625-1679)
and the NVLink broadcast of whole-array inputs on multi-GPU boxes."""
import ctypes as C
import time

import numpy as np
import pytest

import cekirdekler_b200 as ck
from cekirdekler_b200._native import shim

pytestmark = pytest.mark.gpu

SRC = """
__kernel void addOne(__global int* a) { int i = get_global_id(0); a[i] = a[i] + 1; }
__kernel void copyTo(__global int* a, __global int* b) { int i = get_global_id(0); b[i] = a[i]; }
__kernel void spin(__global int* a) { int i = get_global_id(0); int v = a[i]; for (int k = 0; k < 20000; k++) v = v * 3 + 1; a[i] = v; }
__kernel void bump(__global int* counter) { if (get_global_id(0) == 0) counter[0] += 1; }
__kernel void sumNeighbours(__global float* in, __global float* out) { int i = get_global_id(0); out[i] = in[i] + in[(i + 1) % 4096]; }
"""
N = 1 << 16


def dev0():
    return ck.ClPlatforms.all().gpus()[0]


def test_enqueue_mode_defers_synchronisation_and_async_spreads_over_queues(gpus):
    cr = ck.ClNumberCruncher(dev0(), SRC)
    arrays = []
    cr.enqueueMode = True
    for k in range(6):
        a = ck.ClArray(N, np.int32)
        a.view[:] = k
        cr.enqueueModeAsyncEnable = True          # every assignment advances the queue cursor (Cores.cs:83)
        assert a.compute(cr, 100 + k, "addOne", N, 256) == 0
        cr.enqueueModeAsyncEnable = False
        arrays.append(a)
    cr.enqueueMode = False                        # the transition synchronises every queue that was used (Cores.cs:112-119)
    for k, a in enumerate(arrays):
        assert (a.view == k + 1).all()


def test_markers_count_enqueued_and_reached_computes(gpus):
    cr = ck.ClNumberCruncher(dev0(), SRC)
    cr.fineGrainedQueueControl = True
    a = ck.ClArray(N, np.int32)
    a.read, a.write = False, False
    cr.enqueueMode = True
    for _ in range(5):
        cr.enqueueModeAsyncEnable = True
        assert a.compute(cr, 1, "spin", N, 256) == 0
        cr.enqueueModeAsyncEnable = False
    assert cr.countMarkersRemaining() + cr.countMarkersReached() == 5
    cr.enqueueMode = False
    assert cr.countMarkersRemaining() == 0 and cr.countMarkersReached() == 5


def test_no_compute_mode_only_moves_data(gpus):
    cr = ck.ClNumberCruncher(dev0(), SRC)
    a, b = ck.ClArray(N, np.int32), ck.ClArray(N, np.int32)
    a.view[:] = 7
    a.readOnly = True
    b.read = False                                # output only
    assert a.nextParam(b).compute(cr, 1, "copyTo", N, 256) == 0
    assert (b.view == 7).all()
    a.view[:] = 9
    b.view[:] = 0
    cr.noComputeMode = True                       # uploads a, downloads b, launches nothing: b still holds the old copy
    assert a.nextParam(b).compute(cr, 1, "copyTo", N, 256) == 0
    cr.noComputeMode = False
    assert (b.view == 7).all()
    assert a.nextParam(b).compute(cr, 1, "copyTo", N, 256) == 0
    assert (b.view == 9).all()


def test_repeat_count_with_sync_kernel(gpus):
    a, counter = ck.ClArray(N, np.int32), ck.ClArray(256, np.int32)
    a.view[:] = 0
    counter.view[:] = 0
    counter.writeAll = True                       # 256 elements < global range: comes back whole
    g = a.nextParam(counter)
    src2 = SRC + "__kernel void addBoth(__global int* a, __global int* counter) { int i = get_global_id(0); a[i] += 1; }\n" \
                 "__kernel void bump2(__global int* a, __global int* counter) { if (get_global_id(0) == 0) counter[0] += 1; }\n"
    cr2 = ck.ClNumberCruncher(dev0(), src2)
    cr2.repeatCount = 10
    cr2.repeatKernelName = "bump2"                # runs one work-group at offset 0 after every repeat (Cores.cs:793)
    assert g.compute(cr2, 1, "addBoth", N, 256) == 0
    assert (a.view == 10).all() and counter[0] == 10
    cr2.repeatCount = 1
    assert g.compute(cr2, 1, "addBoth", N, 256) == 0
    assert (a.view == 11).all()


def test_two_kernel_repeat_runs_the_sync_kernel(gpus):
    """Several kernels, repeatCount > 1 and a repeatKernelName (Cores.cs:793-800): the sync kernel runs once per repeat
    with the arguments it was bound with."""
    a, counter = ck.ClArray(N, np.int32), ck.ClArray(256, np.int32)
    a.view[:] = 0
    counter.view[:] = 0
    counter.writeAll = True
    src2 = "__kernel void addBoth(__global int* a, __global int* counter) { int i = get_global_id(0); a[i] += 1; }\n" \
           "__kernel void addTwo(__global int* a, __global int* counter) { int i = get_global_id(0); a[i] += 2; }\n" \
           "__kernel void bump2(__global int* a, __global int* counter) { if (get_global_id(0) == 0) counter[0] += 1; }\n"
    cr2 = ck.ClNumberCruncher(dev0(), src2)
    cr2.repeatCount = 5
    cr2.repeatKernelName = "bump2"
    assert a.nextParam(counter).compute(cr2, 1, "addBoth addTwo", N, 256) == 0
    assert (a.view == 15).all() and counter[0] == 5


@pytest.mark.parametrize("ptype", ["event", "driver"])
def test_pipelined_write_all_comes_back_complete(gpus, ptype):
    """A `write all` array of a pipelined compute is fetched whole only after every blob's kernel has finished
    (quirk Q2 fixed, Worker.cs:871-885): a histogram over all blobs must be complete on the host."""
    n = 1 << 22
    src = "__kernel void count(__global const int* v, __global int* hist) { int i = get_global_id(0); atomic_add(&hist[v[i] & 255], 1); }\n" \
          "__kernel void slowCount(__global const int* v, __global int* hist) { int i = get_global_id(0); int x = v[i]; " \
          "for (int k = 0; k < 2000; k++) x = x * 1664525 + 1013904223; atomic_add(&hist[v[i] & 255], (x == 0x12345678 && i == -1) ? 2 : 1); }\n"
    cr = ck.ClNumberCruncher(dev0(), src)
    assert cr.errorCode() == 0, cr.errorMessage()
    v, hist = ck.ClArray(n, np.int32), ck.ClArray(256, np.int32)
    v.view[:] = (np.arange(n, dtype=np.int64) * 2654435761 >> 7).astype(np.int32)
    v.partialRead, v.read, v.write = True, False, False
    hist.read, hist.writeAll = True, True
    exp = np.bincount(v.view & 255, minlength=256)
    for kernel in ("count", "slowCount"):
        hist.view[:] = 0
        assert v.nextParam(hist).compute(cr, 1, kernel, n, 256, 0, True, ck.PIPELINE_EVENT if ptype == "event" else ck.PIPELINE_DRIVER, 8) == 0
        assert np.array_equal(hist.view, exp), kernel


def test_zero_copy_stream_arrays_are_read_in_place(gpus):
    devs = ck.ClPlatforms.all().gpus(streamingEnabled=True)[0]      # GPU_STREAM: device.isGddr() = false (ClDevice.cs:105-108)
    cr = ck.ClNumberCruncher(devs, SRC)
    a, b = ck.ClArray(N, np.int32), ck.ClArray(N, np.int32)
    a.zeroCopy, b.zeroCopy = True, True
    a.view[:] = np.arange(N)
    assert a.nextParam(b).compute(cr, 1, "copyTo", N, 256) == 0
    assert np.array_equal(b.view, np.arange(N))
    a.view[:] = 5                                   # no upload needed: the kernel dereferences the pinned host array itself
    a.read, a.write, b.read, b.write = False, False, False, False
    assert a.nextParam(b).compute(cr, 1, "copyTo", N, 256) == 0
    cr.sync()
    assert (b.view == 5).all()


def test_validation_rules_refuse_and_count_errors(gpus):
    cr = ck.ClNumberCruncher(dev0(), SRC)
    a = ck.ClArray(1000, np.int32)
    assert a.compute(cr, 1, "addOne", 1000, 256) == 3          # global not a multiple of local
    assert cr.numberOfErrorsHappened == 1
    assert a.compute(cr, 1, "addOne", 768, 256) == 1           # any earlier error blocks later computes (ClArray.cs:1610-1616)
    cr = ck.ClNumberCruncher(dev0(), SRC)
    assert a.compute(cr, 1, "addOne", 2048, 256) == 7          # array shorter than global range * elements per item
    cr = ck.ClNumberCruncher(dev0(), SRC)
    a = ck.ClArray(4096, np.int32)
    assert a.compute(cr, 1, "addOne", 4096, 256, 0, True, ck.PIPELINE_EVENT, 3) == 4   # not a multiple of local*blobs
    cr = ck.ClNumberCruncher(dev0(), SRC)
    a3 = ck.ClArray(768, np.int32)
    assert a3.compute(cr, 1, "addOne", 768, 128, 0, True, ck.PIPELINE_EVENT, 3) == 11  # two lanes of blobs/2 chunks: odd counts would skip work
    cr = ck.ClNumberCruncher(dev0(), SRC)
    a3.view[:] = 0
    assert a3.compute(cr, 1, "addOne", 768, 128, 0, True, ck.PIPELINE_DRIVER, 3) == 0  # the driver-controlled pipeline takes any count
    assert (a3.view == 1).all()
    cr = ck.ClNumberCruncher(dev0(), SRC)
    assert a.compute(cr, 1, "noSuchKernel", 4096, 256) == 10   # reported, not raised
    cr = ck.ClNumberCruncher(dev0(), "__kernel void broken(__global int* a) { a[0] = undefined_thing; }")
    assert cr.errorCode() != 0 and "undefined_thing" in cr.errorMessage()
    assert a.compute(cr, 1, "broken", 4096, 256) != 0


def test_performance_report_and_normalised_shares(gpus):
    cr = ck.ClNumberCruncher(ck.ClPlatforms.all().gpus(), SRC)
    a = ck.ClArray(1 << 20, np.int32)
    a.partialRead, a.read = True, False
    for _ in range(4):
        assert a.compute(cr, 5, "addOne", 1 << 20, 256) == 0
    rep = cr.lastComputePerformanceReport(5)
    assert "Compute-ID: 5" in rep and "Load Distributions" in rep and "workitems" in rep
    shares = cr.normalizedGlobalRangesOfDevices(5)
    assert abs(sum(shares) - 1.0) < 1e-9 and len(shares) == gpus
    powers = cr.normalizedComputePowersOfDevices()
    assert abs(sum(powers) - 1.0) < 1e-9
    assert len(cr.deviceNames()) == gpus and all("NVIDIA" in n for n in cr.deviceNames())
    hist = cr.performanceHistory(5)                  # Cores.cs:1075-1098: 10 rows, newest last, one column per device
    assert len(hist) == 10 and all(len(r) == gpus for r in hist)
    assert abs(sum(hist[-1]) - 1.0) < 1e-9 and not any(cr.performanceHistory(77)[0])


def test_whole_read_inputs_are_broadcast_over_nvlink(gpus):
    if gpus < 2:
        pytest.skip("needs two GPUs")
    L = shim()
    n0, p0 = C.c_uint64(), C.c_uint64()
    L.ckBroadcastCounts(C.byref(n0), C.byref(p0))
    cr = ck.ClNumberCruncher(ck.ClPlatforms.all().gpus(), SRC)
    src, out = ck.ClArray(1 << 20, np.float32), ck.ClArray(1 << 20, np.float32)
    src.view[:] = np.arange(1 << 20) % 1000
    src.readOnly = True                            # whole-array read: every device needs all of it
    out.writeOnly = True
    for _ in range(3):
        assert src.nextParam(out).compute(cr, 1, "sumNeighbours", 1 << 20, 256) == 0
    exp = src.view + src.view[(np.arange(1 << 20) + 1) % 4096]
    assert np.array_equal(out.view, exp)
    n1, p1 = C.c_uint64(), C.c_uint64()
    L.ckBroadcastCounts(C.byref(n1), C.byref(p1))
    assert (n1.value - n0.value) + (p1.value - p0.value) == 3      # one device-to-device fan-out per call, not D host uploads
    assert n1.value - n0.value == 3, "NCCL was expected to carry the broadcast on this box"
