"""SingleGPUPipeline.DevicePipeline (ClPipeline.cs:2357-3240): kernel stages overlapped on one device with INPUT / OUTPUT /
INTERNAL / TRANSITION arrays, in serial and in parallel mode.  Checked against plain numpy arithmetic (integer data: exact)."""
import numpy as np
import pytest

import cekirdekler_b200 as ck
from cekirdekler_b200.singlegpupipeline import (DevicePipeline, DevicePipelineArray, DevicePipelineArrayType as T,
                                                DevicePipelineStage)

pytestmark = pytest.mark.gpu

SRC = """
__kernel void addBias(__global int* in, __global int* bias, __global int* t1) { int i = get_global_id(0); t1[i] = in[i] + bias[i]; }
__kernel void timesThree(__global int* t1, __global int* t2)                   { int i = get_global_id(0); t2[i] = t1[i] * 3; }
__kernel void minusOne(__global int* t2, __global int* out)                    { int i = get_global_id(0); out[i] = t2[i] - 1; }
__kernel void setBias(__global int* bias)                                      { int i = get_global_id(0); bias[i] = i % 7; }
"""
N = 1 << 16


def build(serial):
    dev = ck.ClPlatforms.all().gpus()[0]
    inp, out = ck.ClArray(N, np.int32), ck.ClArray(N, np.int32)
    a_in, a_out = DevicePipelineArray(T.INPUT, inp), DevicePipelineArray(T.OUTPUT, out)
    bias = DevicePipelineArray(T.INTERNAL, (np.int32, N))
    t1, t2 = DevicePipelineArray(T.TRANSITION, (np.int32, N)), DevicePipelineArray(T.TRANSITION, (np.int32, N))
    s0 = DevicePipelineStage("addBias", N, 256)
    for a in (a_in, bias, t1):
        s0.bindArray(a)
    s1 = DevicePipelineStage("timesThree", N, 256)
    for a in (t1, t2):
        s1.bindArray(a)
    s2 = DevicePipelineStage("minusOne", N, 256)
    for a in (t2, a_out):
        s2.bindArray(a)
    p = DevicePipeline(dev, SRC)
    for s in (s0, s1, s2):
        p.addStage(s)
    if serial:
        p.enableSerialMode()
    else:
        p.enableParallelMode()
    return p, inp, out, (s0, s1, s2, a_in, a_out, bias, t1, t2)


def expected(k):
    # bias stays zero (INTERNAL arrays start zeroed and nothing writes them in these stages)
    return ((np.arange(N, dtype=np.int64) % 1000 + k) * 3 - 1).astype(np.int32)


@pytest.mark.parametrize("serial", [True, False], ids=["serial", "parallel"])
def test_three_stage_device_pipeline(gpus, serial):
    p, inp, out, keep = build(serial)
    lat = p.latency
    assert lat == (0 if serial else 2)
    for k in range(8):
        inp.view[:] = np.arange(N) % 1000 + k
        p.feed()
        assert p.errorCode() == 0, p.errorMessage()
        if k >= lat:
            assert np.array_equal(out.view, expected(k - lat)), f"feed {k}"


def test_feed_async_runs_host_work_while_the_device_computes(gpus):
    p, inp, out, keep = build(serial=False)
    ran = []
    for k in range(5):
        inp.view[:] = np.arange(N) % 1000 + k
        p.feedAsync(lambda: ran.append(k))
    assert ran == [0, 1, 2, 3, 4]
    assert np.array_equal(out.view, expected(4 - p.latency))


def test_stop_host_device_transmission_keeps_data_resident(gpus):
    p, inp, out, keep = build(serial=True)
    inp.view[:] = np.arange(N) % 1000
    p.feed()
    first = out.view.copy()
    keep[0].stopHostDeviceTransmission = True      # stage 0 stops uploading: the device keeps the old input
    inp.view[:] = 12345
    p.feed()
    assert np.array_equal(out.view, first)
