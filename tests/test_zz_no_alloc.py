"""ckNBody3DStepNoAlloc — the n-body step for callers that cannot allocate device memory.

Every other path in this library hands a kernel its memory through createBuffer, i.e. through cudaMalloc.  That is fine at
set-up time, but cudaMalloc drains every queue of the GPU and can fail: code that runs while an enqueue-mode region, a
pipeline or a device pool has the GPU's queues full, and hosts whose device memory is owned by another allocator, need a
way to run the one kernel here that wants scratch memory (nBody3Dw: unit queue + partial sums) without it.  The entry
point keeps that scratch in a fixed 16 MiB array inside the library's device image and takes plain device pointers.

This file is named so that pytest collects it last: if it fails, `pytest -x` has already run everything else.  The test
drives both paths through the C ABI only — the new entry point directly, and the existing one (`compute` on an
NVRTC-built nBody3Dw with a caller-made scratch buffer) through the host layer — and compares them after each of two
consecutive calls, with the tolerance the other n-body tests use (1e-4 relative on positions), and also checks both calls
against the CPU oracle directly.  A second test takes the entry point to the sizes where its fixed scratch starts to
matter: a share large enough that the kernel has to widen its chunks of partial sums, and the largest share it accepts.
The tests depend on nothing the other tests leave behind, so they pass alone in a fresh process as well as after them."""
import ctypes as C

import numpy as np
import pytest

import cekirdekler_b200 as ck
from cekirdekler_b200 import kernels
from cekirdekler_b200._native import shim
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu

N = 16384
DT, EPS2 = 1e-3, 1e-4


def inputs():
    pos = np.zeros((N, 4), np.float32)
    for c in range(3):
        pos[:, c] = po.uniform(0xC0FFEE + 20 + c, N, -15, 15)
    pos[:, 3] = 1.0
    return pos


def existing_entry_point(pos0, steps):
    """The array-backed kernel, one device, positions after each step."""
    cr = ck.ClNumberCruncher(ck.ClPlatforms.all().gpus()[0], kernels.source("nbody"))
    assert cr.errorCode() == 0, cr.errorMessage()
    kfloats = ((N // 128 + 8) * 8 * 384 + 4096) // 1024
    pa, pb, vel, params, scratch = (ck.ClFloatArray(4 * N), ck.ClFloatArray(4 * N), ck.ClFloatArray(4 * N), ck.ClFloatArray(8),
                                    ck.ClFloatArray(kfloats * 1024))
    pa.view[:] = pos0.reshape(-1)
    params.view[:5] = [DT, EPS2, N, 4, kfloats]
    params.readOnly = True
    scratch.read, scratch.write = False, False
    vel.partialRead, vel.read = True, False
    for a in (pa, pb, vel):
        a.numberOfElementsPerWorkItem = 1
    out, src, dst = [], pa, pb
    for _ in range(steps):
        src.read, src.partialRead, src.write = True, False, False
        dst.read, dst.partialRead, dst.write = False, False, True
        assert src.nextParam(vel, dst, params, scratch).compute(cr, 1, "nBody3Dw", 4 * N, 128) == 0
        out.append((dst.view.reshape(N, 4).copy(), vel.view.reshape(N, 4).copy()))
        src, dst = dst, src
    for a in (pa, pb, vel, params, scratch):
        a.dispose()
    cr.dispose()
    return out


class RawDevice:
    """Context, one queue and float buffers made with the reference's own ABI calls (the set-up a caller does up front)."""

    def __init__(self, index=0):
        L = self.L = shim()
        self.plist = L.platformList()
        self.plat = L.createPlatform(self.plist, 0)
        self.dev = L.createDevice(self.plat, L.CODE_GPU(), index)
        self.ctx = L.createContext(self.plat, self.dev)
        self.queue = L.createCommandQueue(self.ctx, self.dev, 0)
        assert self.queue, L.ckGetLastError()
        self.ordinal = L.ckContextDeviceOrdinal(self.ctx)
        self.stream = L.ckQueueStream(self.queue)
        self.buffers = []

    def buffer(self, host: np.ndarray):
        b = self.L.createBuffer(self.ctx, host.size, 0, 0, host.ctypes.data_as(C.c_void_p), 1, 0, 0)
        assert b, self.L.ckGetLastError()
        self.buffers.append(b)
        return b

    def close(self):
        L = self.L
        L.finish(self.queue)
        for b in self.buffers:
            L.deleteBuffer(b)
        L.deleteCommandQueue(self.queue)
        L.deleteContext(self.ctx)
        L.deleteDevice(self.dev)
        L.deletePlatform(self.plat)
        L.deletePlatformList(self.plist)


def test_no_alloc_step_matches_the_array_backed_kernel_on_two_calls_in_a_row(gpus):
    L = shim()
    pos0 = inputs()
    expected = existing_entry_point(pos0, 2)

    d = RawDevice()
    try:
        h_a, h_b, h_v = pos0.reshape(-1).copy(), np.zeros(4 * N, np.float32), np.zeros(4 * N, np.float32)
        b_a, b_b, b_v = d.buffer(h_a), d.buffer(h_b), d.buffer(h_v)
        L.writeToBuffer(d.queue, b_a, h_a.ctypes.data_as(C.c_void_p))
        L.writeToBuffer(d.queue, b_v, h_v.ctypes.data_as(C.c_void_p))
        p_a, p_b, p_v = (L.ckBufferDevicePointer(b) for b in (b_a, b_b, b_v))
        launches = L.ckLaunchCount()
        o_p, o_v, o_o = pos0.reshape(-1).copy(), np.zeros(4 * N, np.float32), np.zeros(4 * N, np.float32)
        src, dst, h_dst = p_a, p_b, h_b
        b_dst = b_b
        for step in range(2):
            # two half-device calls per step, like two shares of a split; same stream, so they do not overlap
            for first, count in ((0, N // 2), (N // 2, N // 2)):
                rc = L.ckNBody3DStepNoAlloc(d.ordinal, d.stream, src, p_v, dst, N, first, count, DT, EPS2)
                assert rc == 0, L.ckGetLastError()
            L.readFromBuffer(d.queue, b_dst, h_dst.ctypes.data_as(C.c_void_p))
            L.readFromBuffer(d.queue, b_v, h_v.ctypes.data_as(C.c_void_p))
            L.finish(d.queue)
            exp_pos, exp_vel = expected[step]
            got = h_dst.reshape(N, 4)
            rel = np.abs(got[:, :3] - exp_pos[:, :3]) / np.maximum(np.abs(exp_pos[:, :3]), 1e-3)
            assert rel.max() <= 1e-4, (step, rel.max())
            assert np.array_equal(got[:, 3], exp_pos[:, 3])
            vrel = np.abs(h_v.reshape(N, 4)[:, :3] - exp_vel[:, :3]).max() / np.abs(exp_vel[:, :3]).max()
            assert vrel <= 1e-4, (step, vrel)
            # and against the CPU oracle itself, not only against the library's other GPU path
            po.nbody3d_step(o_p, o_v, o_o, np.array([DT, EPS2, N, 0], np.float32))
            o_p, o_o = o_o, o_p
            oexp = o_p.reshape(N, 4)
            orel = np.abs(got[:, :3] - oexp[:, :3]) / np.maximum(np.abs(oexp[:, :3]), 1e-3)
            assert orel.max() <= 1e-4, (step, orel.max())
            assert np.array_equal(got[:, 3], oexp[:, 3])
            src, dst = dst, src
            h_dst, b_dst = (h_a, b_a) if h_dst is h_b else (h_b, b_b)
        assert L.ckLaunchCount() == launches          # none of the NVRTC path's kernels ran for these calls

        # beyond the stated limits, or malformed: refused with a message, nothing launched, later calls unaffected
        assert L.ckNBody3DStepNoAlloc(d.ordinal, d.stream, p_a, p_v, p_b, 1 << 25, 0, 1 << 21, DT, EPS2) == 2
        assert b"exceed" in L.ckGetLastError()
        assert L.ckNBody3DStepNoAlloc(d.ordinal, d.stream, p_a, p_v, p_b, N, 0, 48, DT, EPS2) == 1
        assert L.ckNBody3DStepNoAlloc(d.ordinal, d.stream, None, p_v, p_b, N, 0, N, DT, EPS2) == 1
        assert L.ckNBody3DStepNoAlloc(d.ordinal, d.stream, p_a, p_v, p_b, N, N - 32, 64, DT, EPS2) == 1
        assert L.ckNBody3DStepNoAlloc(d.ordinal, d.stream, src, p_v, dst, N, 0, N, DT, EPS2) == 0
        L.finish(d.queue)
    finally:
        d.close()


@pytest.mark.parametrize("n,first,bodies,samples", [(131072, 4096, 32768, 1024), (1048576, 0, 1048576, 256)])
def test_no_alloc_step_at_the_sizes_where_the_fixed_scratch_matters(gpus, n, first, bodies, samples):
    """32,768 bodies of 131,072 in one call need more partial sums than 16 MiB hold at the kernel's preferred chunking, so the
    kernel widens its chunks (different summation order, same 1e-4 tolerance); 1,048,576 bodies is the most one call accepts.
    Both are checked against the CPU oracle on bodies spread over the share (the oracle takes an offset and a count), and
    the bodies outside the share must be left alone."""
    L = shim()
    pos = np.zeros((n, 4), np.float32)
    for c in range(3):
        pos[:, c] = po.uniform(0xC0FFEE + 30 + c, n, -15, 15)
    pos[:, 3] = 1.0
    d = RawDevice()
    try:
        h_p, h_o, h_v = pos.reshape(-1).copy(), np.full(4 * n, -7.0, np.float32), np.zeros(4 * n, np.float32)
        b_p, b_o, b_v = d.buffer(h_p), d.buffer(h_o), d.buffer(h_v)
        for b, h in ((b_p, h_p), (b_o, h_o), (b_v, h_v)):
            L.writeToBuffer(d.queue, b, h.ctypes.data_as(C.c_void_p))
        p_p, p_o, p_v = (L.ckBufferDevicePointer(b) for b in (b_p, b_o, b_v))
        rc = L.ckNBody3DStepNoAlloc(d.ordinal, d.stream, p_p, p_v, p_o, n, first, bodies, DT, EPS2)
        assert rc == 0, L.ckGetLastError()
        L.readFromBuffer(d.queue, b_o, h_o.ctypes.data_as(C.c_void_p))
        L.readFromBuffer(d.queue, b_v, h_v.ctypes.data_as(C.c_void_p))
        L.finish(d.queue)
        got, gotv = h_o.reshape(n, 4), h_v.reshape(n, 4)
        outside = np.ones(n, bool)
        outside[first:first + bodies] = False
        assert (got[outside] == -7.0).all() and not gotv[outside].any()
        params = np.array([DT, EPS2, n, 0], np.float32)
        block = 64
        for k in range(samples // block):
            f = first + (k * (bodies - block)) // max(1, samples // block - 1)
            v, o = np.zeros(4 * n, np.float32), np.zeros(4 * n, np.float32)
            po.nbody3d_step(h_p, v, o, params, f, block)
            exp = o.reshape(n, 4)[f:f + block]
            rel = np.abs(got[f:f + block, :3] - exp[:, :3]) / np.maximum(np.abs(exp[:, :3]), 1e-3)
            assert rel.max() <= 1e-4, (f, rel.max())
            assert np.array_equal(got[f:f + block, 3], exp[:, 3])
            vexp = v.reshape(n, 4)[f:f + block, :3]
            assert np.abs(gotv[f:f + block, :3] - vexp).max() <= 1e-4 * np.abs(vexp).max(), f
    finally:
        d.close()


def test_no_alloc_step_on_a_second_device_has_its_own_scratch(gpus):
    """Each GPU holds its own copy of the built-in array and the entry point looks its address up per device: a call on GPU 1
    right after one on GPU 0 must give the same bits as GPU 0 did."""
    if gpus < 2:
        pytest.skip("needs two GPUs")
    L = shim()
    pos0 = inputs()
    results = []
    for index in (0, 1, 0):
        d = RawDevice(index)
        try:
            h_p, h_o, h_v = pos0.reshape(-1).copy(), np.zeros(4 * N, np.float32), np.zeros(4 * N, np.float32)
            b_p, b_o, b_v = d.buffer(h_p), d.buffer(h_o), d.buffer(h_v)
            L.writeToBuffer(d.queue, b_p, h_p.ctypes.data_as(C.c_void_p))
            L.writeToBuffer(d.queue, b_v, h_v.ctypes.data_as(C.c_void_p))
            p_p, p_o, p_v = (L.ckBufferDevicePointer(b) for b in (b_p, b_o, b_v))
            assert L.ckNBody3DStepNoAlloc(d.ordinal, d.stream, p_p, p_v, p_o, N, 0, N, DT, EPS2) == 0, L.ckGetLastError()
            L.readFromBuffer(d.queue, b_o, h_o.ctypes.data_as(C.c_void_p))
            L.finish(d.queue)
            results.append(h_o.copy())
        finally:
            d.close()
    assert results[0].any() and np.array_equal(results[0], results[1]) and np.array_equal(results[0], results[2])
