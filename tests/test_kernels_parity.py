"""CUDA kernels, launched through ClNumberCruncher → libcekirdekler_host.so → the KutuphaneCL C ABI, against the CPU
oracle on the same seeded inputs and against the committed golden fixtures.
Bars: bit-exact for Mandelbrot iteration counts, vector-add, SAXPY (contraction off), 2-D n-body forces, blur, Sobel and
the histogram; <= 1e-4 relative on 3-D n-body positions (the GPU uses the SFU reciprocal square root)."""
import hashlib
import json
from pathlib import Path

import numpy as np
import pytest

import cekirdekler_b200 as ck
import reference_kernels
from cekirdekler_b200 import kernels
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu
GOLDEN = json.loads((Path(__file__).resolve().parent / "golden" / "kernel_vectors.json").read_text())


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def gpu0():
    return ck.ClPlatforms.all().gpus()[0]


def all_gpus():
    return ck.ClPlatforms.all().gpus()


# ------------------------------------------------------------------ Mandelbrot (K1)
def run_mandelbrot(devs, W, H, max_iter, kernel="mandelbrot", calls=1, pipeline=False, pipeline_type=True, blobs=4):
    cr = ck.ClNumberCruncher(devs, kernels.source("mandelbrot"))
    assert cr.errorCode() == 0, cr.errorMessage()
    view = ck.ClArray(8, np.float32)
    view.view[:] = po.mandelbrot_view(W, H, max_iter)
    view.readOnly = True                      # whole-array `read` input (scalars travel in arrays: F9)
    out = ck.ClArray(W * H, np.int32)
    out.writeOnly = True                      # sliced `write` output
    e = 4 if kernel == "mandelbrot" else 1
    out.numberOfElementsPerWorkItem = e
    for _ in range(calls):
        out.view[:] = -1
        rc = view.nextParam(out).compute(cr, 1, kernel, W * H // e, 256, 0, pipeline, pipeline_type, blobs)
        assert rc == 0
    return out.view.copy(), cr


def test_mandelbrot_c1_bit_exact_against_oracle_and_golden(gpus):
    got, _ = run_mandelbrot(gpu0(), 1024, 1024, 256)
    exp = po.mandelbrot(po.mandelbrot_view(1024, 1024, 256), 1024 * 1024)
    assert np.array_equal(got, exp)
    g = GOLDEN["mandelbrot_c1"]
    assert sha(got) == g["sha256"] and int(got.sum()) == g["sum"]
    crop = np.array(g["crop_rows_480_544_cols_200_264"], np.int32)
    assert np.array_equal(got.reshape(1024, 1024)[480:544, 200:264], crop)


@pytest.mark.parametrize("kernel", ["mandelbrot", "mandelbrot1"])
@pytest.mark.parametrize("max_iter", [1, 7, 8, 9, 31, 32, 33, 95, 96, 97, 100, 1000])
def test_mandelbrot_iteration_limits(gpus, kernel, max_iter):
    # max_iter around the unchecked-block boundaries (4 blocks of 8, then blocks of 64) exercises the block/tail/replay boundaries
    got, _ = run_mandelbrot(gpu0(), 512, 256, max_iter, kernel)
    exp = po.mandelbrot(po.mandelbrot_view(512, 256, max_iter), 512 * 256)
    assert np.array_equal(got, exp)


def test_mandelbrot_zoomed_views_bit_exact(gpus):
    cr = ck.ClNumberCruncher(gpu0(), kernels.source("mandelbrot"))
    W = H = 512
    out = ck.ClArray(W * H, np.int32)
    out.writeOnly = True
    out.numberOfElementsPerWorkItem = 4
    view = ck.ClArray(8, np.float32)
    view.readOnly = True
    for (x0, x1, y0, y1, it) in [(-2.0, -1.9, -0.05, 0.05, 2000), (-0.75, -0.74, 0.1, 0.11, 3000),
                                 (-2.5, 2.5, -2.5, 2.5, 64), (0.25, 0.26, 0.0, 0.01, 4096)]:
        v = po.mandelbrot_view(W, H, it, x0, x1, y0, y1)
        view.view[:] = v
        assert view.nextParam(out).compute(cr, 7, "mandelbrot", W * H // 4, 256) == 0
        assert np.array_equal(out.view, po.mandelbrot(v, W * H)), (x0, x1, y0, y1)


def test_mandelbrot_c3_strided_sample_matches_golden(gpus):
    """Full C3 size (16384^2 x 4096) on the GPU; checked on the committed strided oracle sample and by the
    size-independent property that the image is symmetric about the real axis (rows y and H-1-y... the view is
    symmetric: cy(py) = -cy(H-py) exactly only for py and H-py, so rows 1..H-1 mirror)."""
    g = GOLDEN["mandelbrot_c3_sample"]
    W = H = 16384
    got, _ = run_mandelbrot(gpu0(), W, H, 4096)
    idx = g["offset"] + g["stride"] * np.arange(g["count"], dtype=np.int64)
    sample = got[idx]
    assert sha(sample.astype(np.int32)) == g["sha256"] and int(sample.sum()) == g["sum"]
    img = got.reshape(H, W)
    # cy = fma(py, dy, y0) with y0 = -1.5, dy = 3/16384 (exact in binary): cy(py) == -cy(H - py) bit-for-bit, and the
    # iteration is symmetric under conjugation, so row py equals row H - py for py in [1, H-1]
    assert np.array_equal(img[1:H // 2], img[H - 1:H // 2:-1])


def test_mandelbrot_multi_gpu_balanced_and_pipelined(gpus):
    W, H, it = 2048, 2048, 512
    exp = po.mandelbrot(po.mandelbrot_view(W, H, it), W * H)
    got, cr = run_mandelbrot(all_gpus(), W, H, it, calls=6)
    assert np.array_equal(got, exp)
    r = cr.globalRanges(1)
    assert sum(r) == W * H // 4 and all(x % 256 == 0 for x in r)
    for ptype in (True, False):
        got, _ = run_mandelbrot(all_gpus(), W, H, it, calls=3, pipeline=True, pipeline_type=ptype, blobs=8)
        assert np.array_equal(got, exp)


@pytest.mark.parametrize("W,H", [(1024, 1024), (256, 64), (512, 23), (96, 96), (2048, 2048)])
def test_mandelbrot_one_pixel_kernel_bit_exact(gpus, W, H):
    """mandelbrot1 (the benchmark's kernel): same counts as the oracle on one device and on all, plain and pipelined."""
    it = 300
    exp = po.mandelbrot(po.mandelbrot_view(W, H, it), W * H)
    for devs in (gpu0(), all_gpus()):
        got, cr = run_mandelbrot(devs, W, H, it, "mandelbrot1", calls=3)
        assert np.array_equal(got, exp)
        if W * H % (256 * 4) == 0 and W * H >= 256 * 4 * len(cr.globalRanges(1)):   # the pipelines' own size rule (ClArray.cs:1640-1660)
            for ptype in (True, False):
                got, _ = run_mandelbrot(devs, W, H, it, "mandelbrot1", calls=2, pipeline=True, pipeline_type=ptype, blobs=4)
                assert np.array_equal(got, exp)


def test_mandelbrot_with_a_range_offset(gpus):
    """A compute() that starts in the middle of the image (global offset) and covers a ragged number of rows."""
    W, H, it = 512, 64, 200
    v = po.mandelbrot_view(W, H, it)
    exp = po.mandelbrot(v, W * H)
    cr = ck.ClNumberCruncher(gpu0(), kernels.source("mandelbrot"))
    view, out = ck.ClArray(8, np.float32), ck.ClArray(W * H, np.int32)
    view.view[:] = v
    view.readOnly = True
    out.writeOnly = True
    out.view[:] = -1
    first, count = 3 * W + 256, 41 * W + 256
    assert view.nextParam(out).compute(cr, 1, "mandelbrot1", count, 256, first) == 0
    assert np.array_equal(out.view[first:first + count], exp[first:first + count])
    assert (out.view[:first] == -1).all() and (out.view[first + count:] == -1).all()


# ------------------------------------------------------------------ vector-add / SAXPY (K2, K3)
def test_reference_vector_add_test(gpus):
    """Tester.stream_C_equals_A_plus_B_1M_elements (Tester.cs:7806-7843): API style 2 flag strings, driver pipeline with
    8 blobs, local 256, mixed array kinds (native array, language array, ClArray), 10 calls; the reference checks c[700]."""
    n = 1024 * 1024
    a = ck.ClFloatArray(n)
    b_host = np.arange(n, dtype=np.float32)
    b = ck.ClArray(b_host)
    c = ck.ClArray(n, np.float32)
    a.view[:] = 3
    c.view[:] = 105
    cr = ck.ClNumberCruncher(all_gpus(), reference_kernels.REFERENCE_VECTOR_ADD)
    assert cr.errorCode() == 0, cr.errorMessage()
    g = a.nextParam(b, c)
    for _ in range(10):
        rc = g.compute(cr, 1, "vectorAdd", n, 256, 0, True, ck.PIPELINE_DRIVER, 8,
                       readWritesReady=["partial read", "partial read", "write"], elementsPerItemReady=[1, 1, 1])
        assert rc == 0
    assert abs(c[700] - (a[700] + b[700])) < 0.5       # the reference's own check
    exp = np.zeros(n, np.float32)
    po.vector_add(a.view, b_host, exp)
    assert np.array_equal(c.view, exp)                  # and bit-exact everywhere


@pytest.mark.parametrize("mode", ["plain", "event", "driver"])
def test_vector_add4_and_saxpy4_bit_exact(gpus, mode):
    n = 1 << 22
    cr = ck.ClNumberCruncher(all_gpus(), kernels.source("stream"))
    assert cr.errorCode() == 0, cr.errorMessage()
    a, b, c = ck.ClFloatArray(n), ck.ClFloatArray(n), ck.ClFloatArray(n)
    a.view[:] = po.uniform(1, n, -100, 100)
    b.view[:] = po.uniform(2, n, -100, 100)
    for arr in (a, b):
        arr.partialRead, arr.read, arr.readOnly = True, False, True
        arr.numberOfElementsPerWorkItem = 4
    c.writeOnly = True
    c.numberOfElementsPerWorkItem = 4
    pipe = mode != "plain"
    for _ in range(3):
        assert a.nextParam(b, c).compute(cr, 1, "vectorAdd4", n // 4, 256, 0, pipe, mode == "event", 16) == 0
    exp = np.zeros(n, np.float32)
    po.vector_add(a.view, b.view, exp)
    assert np.array_equal(c.view, exp)

    alpha = ck.ClFloatArray(1)
    alpha[0] = 2.0
    alpha.readOnly = True
    x, y = a, ck.ClFloatArray(n)
    y.view[:] = b.view
    y.partialRead, y.read = True, False
    y.numberOfElementsPerWorkItem = 4
    exp = b.view.copy()
    for _ in range(2):
        assert alpha.nextParam(x, y).compute(cr, 2, "saxpy4", n // 4, 256, 0, pipe, mode == "event", 16) == 0
        po.saxpy(alpha.view, x.view, exp)
    assert np.array_equal(y.view, exp)


def test_saxpy_scalar_kernel_and_exactly_representable_inputs(gpus):
    # BASELINE's C2 inputs: a[i] = 3, b[i] = i mod 2^20, alpha = 2 → every sum is exact
    n = 1 << 20
    cr = ck.ClNumberCruncher(gpu0(), kernels.source("stream"))
    alpha, x, y = ck.ClFloatArray(1), ck.ClFloatArray(n), ck.ClFloatArray(n)
    alpha[0] = 2.0
    x.view[:] = 3
    y.view[:] = np.arange(n) % (1 << 20)
    alpha.readOnly = True
    x.readOnly = True
    assert alpha.nextParam(x, y).compute(cr, 1, "saxpy", n, 256) == 0
    assert np.array_equal(y.view, (6 + np.arange(n) % (1 << 20)).astype(np.float32))


# ------------------------------------------------------------------ n-body (K4)
def test_nbody2d_reference_kernel_text_and_tiled_kernel(gpus):
    """Tester.nBody (Tester.cs:7682-7799) at n = 8192, local 64: the reference's own kernel text (n spliced in) and the
    shared-memory-tiled restatement, both against the reference's host loop (oracle_nbody2d).  The reference accepts
    ±0.01 absolute.  Its kernel text leaves the compiler free to fuse multiply-adds (as an OpenCL compiler is), so it is
    held to the reference's own tolerance; the restatement uses explicit non-fused operations and is bit-exact."""
    n = 8192
    x, y = po.uniform(0xC0FFEE, n, -15, 15), po.uniform(0xC0FFEE + 1, n, -15, 15)
    fx_exp, fy_exp = np.zeros(n, np.float32), np.zeros(n, np.float32)
    po.nbody2d(x, y, fx_exp, fy_exp)

    cr = ck.ClNumberCruncher(all_gpus(), reference_kernels.reference_nbody_kernel(n))
    assert cr.errorCode() == 0, cr.errorMessage()
    fx, fy = np.zeros(n, np.float32), np.zeros(n, np.float32)
    xc, yc, fxc, fyc = ck.ClArray(x), ck.ClArray(y), ck.ClArray(fx), ck.ClArray(fy)
    for _ in range(5):
        assert xc.nextParam(yc, fxc, fyc).compute(cr, 1, "nBody", n, 64) == 0
    assert np.all(np.abs(fx - fx_exp) <= 0.01) and np.all(np.abs(fy - fy_exp) <= 0.01)   # the reference's check
    assert np.allclose(fx, fx_exp, rtol=1e-4, atol=1e-4) and np.array_equal(fx, fy)

    cr2 = ck.ClNumberCruncher(all_gpus(), kernels.source("nbody"))
    assert cr2.errorCode() == 0, cr2.errorMessage()
    fx2, fy2 = ck.ClFloatArray(n), ck.ClFloatArray(n)
    nn = ck.ClIntArray(1)
    nn[0] = n
    nn.readOnly = True
    xc.readOnly, yc.readOnly = True, True
    fx2.writeOnly, fy2.writeOnly = True, True
    assert xc.nextParam(yc, fx2, fy2, nn).compute(cr2, 1, "nBody2D", n, 256) == 0
    assert np.array_equal(fx2.view, fx_exp) and np.array_equal(fy2.view, fy_exp)


def test_nbody2d_golden(gpus):
    n = 1024
    g = GOLDEN["nbody2d_1024"]
    x, y = po.uniform(g["seed_x"], n, -15, 15), po.uniform(g["seed_y"], n, -15, 15)
    cr = ck.ClNumberCruncher(gpu0(), kernels.source("nbody"))
    fx, fy, nn = ck.ClFloatArray(n), ck.ClFloatArray(n), ck.ClIntArray(1)
    nn[0] = n
    nn.readOnly = True
    xc, yc = ck.ClArray(x), ck.ClArray(y)
    assert xc.nextParam(yc, fx, fy, nn).compute(cr, 1, "nBody2D", n, 256) == 0
    assert sha(fx.view) == g["sha256_fx"]
    assert np.array_equal(fx.view[:8], np.array(g["fx_head"], np.float32))


def nbody3d_inputs(n, seed=0xC0FFEE + 10):
    pos = np.zeros((n, 4), np.float32)
    for c in range(3):
        pos[:, c] = po.uniform(seed + c, n, -15, 15)
    pos[:, 3] = 1.0
    return pos


@pytest.mark.parametrize("kernel,items_per_body,local", [("nBody3D", 1, 128), ("nBody3D", 1, 256), ("nBody3D1", 1, 256), ("nBody3Dq", 4, 256)])
@pytest.mark.parametrize("n,steps", [(2048, 1), (16384, 10)])
def test_nbody3d_positions_within_1e4_relative(gpus, n, steps, kernel, items_per_body, local):
    pos0 = nbody3d_inputs(n)
    params_h = np.array([1e-3, 1e-4, n, 0], np.float32)
    # oracle
    p, v, o = pos0.copy().reshape(-1), np.zeros(4 * n, np.float32), np.zeros(4 * n, np.float32)
    for _ in range(steps):
        po.nbody3d_step(p, v, o, params_h)
        p, o = o, p
    # GPU: pos is a whole-array read (broadcast to every device), vel/posOut are sliced 4 floats per work-item
    cr = ck.ClNumberCruncher(all_gpus(), kernels.source("nbody"))
    assert cr.errorCode() == 0, cr.errorMessage()
    pa, pb, vel, params = ck.ClFloatArray(4 * n), ck.ClFloatArray(4 * n), ck.ClFloatArray(4 * n), ck.ClFloatArray(4)
    pa.view[:] = pos0.reshape(-1)
    params.view[:] = params_h
    params.readOnly = True
    vel.partialRead, vel.read = True, False
    vel.numberOfElementsPerWorkItem = 4 // items_per_body
    for arr in (pa, pb):
        arr.numberOfElementsPerWorkItem = 4 // items_per_body
    src, dst = pa, pb
    for _ in range(steps):
        src.read, src.partialRead, src.write = True, False, False      # input: whole array to every device
        dst.read, dst.partialRead, dst.write = False, False, True      # output: each device's slice
        assert src.nextParam(vel, dst, params).compute(cr, 1, kernel, n * items_per_body, local) == 0
        src, dst = dst, src
    got = src.view.reshape(n, 4)
    exp = p.reshape(n, 4)
    rel = np.abs(got[:, :3] - exp[:, :3]) / np.maximum(np.abs(exp[:, :3]), 1e-3)
    assert rel.max() <= 1e-4, rel.max()
    assert np.array_equal(got[:, 3], exp[:, 3])
    if n == 2048:
        g = GOLDEN["nbody3d_2048"]
        ref = np.array(g["pos_out_head"], np.float32)
        assert np.allclose(got[:4], ref, rtol=1e-4, atol=1e-6)


def run_nbody3dw(devs, pos0, steps, wpb, local, scratch_kfloats=None, balance_steps=0):
    """nBody3Dw through compute(): scratch is a device-resident array (no read, no write); returns positions and velocities."""
    n = pos0.shape[0]
    cr = ck.ClNumberCruncher(devs, kernels.source("nbody"))
    assert cr.errorCode() == 0, cr.errorMessage()
    groups, chunks = (n + 127) // 128 + 8, ((n + 255) // 256 + 2 * (local // 32) - 1) // (2 * (local // 32))
    kfloats = scratch_kfloats if scratch_kfloats is not None else (groups * chunks * 384 + groups + 256 + 1023) // 1024
    pa, pb, vel, params, scratch = (ck.ClFloatArray(4 * n), ck.ClFloatArray(4 * n), ck.ClFloatArray(4 * n), ck.ClFloatArray(8),
                                    ck.ClFloatArray(kfloats * 1024))
    pa.view[:] = pos0.reshape(-1)
    params.view[:5] = [1e-3, 1e-4, n, wpb, kfloats]
    params.readOnly = True
    scratch.read, scratch.write = False, False
    vel.partialRead, vel.read = True, False
    for arr in (pa, pb, vel):
        arr.numberOfElementsPerWorkItem = 4 // wpb
    src, dst = pa, pb
    for _ in range(steps):
        src.read, src.partialRead, src.write = True, False, False
        dst.read, dst.partialRead, dst.write = False, False, True
        assert src.nextParam(vel, dst, params, scratch).compute(cr, 1, "nBody3Dw", n * wpb, local) == 0
        src, dst = dst, src
    out = src.view.reshape(n, 4).copy(), vel.view.reshape(n, 4).copy()
    for a in (pa, pb, vel, params, scratch):
        a.dispose()
    cr.dispose()
    return out


@pytest.mark.parametrize("wpb,local", [(1, 128), (4, 128), (4, 256), (2, 64), (4, 32)])
@pytest.mark.parametrize("n,steps", [(2048, 1), (16384, 10), (8192 + 512, 3)])
def test_nbody3d_work_queue_positions_within_1e4_relative(gpus, n, steps, wpb, local):
    """Ten steps in a row also prove that the kernel re-arms its own queue; 8704 bodies leave ragged groups and a short last chunk."""
    pos0 = nbody3d_inputs(n)
    params_h = np.array([1e-3, 1e-4, n, 0], np.float32)
    p, v, o = pos0.copy().reshape(-1), np.zeros(4 * n, np.float32), np.zeros(4 * n, np.float32)
    for _ in range(steps):
        po.nbody3d_step(p, v, o, params_h)
        p, o = o, p
    got, vel = run_nbody3dw(all_gpus(), pos0, steps, wpb, local)
    exp = p.reshape(n, 4)
    rel = np.abs(got[:, :3] - exp[:, :3]) / np.maximum(np.abs(exp[:, :3]), 1e-3)
    assert rel.max() <= 1e-4, rel.max()
    assert np.array_equal(got[:, 3], exp[:, 3])
    vrel = np.abs(vel[:, :3] - v.reshape(n, 4)[:, :3]).max() / np.abs(v.reshape(n, 4)[:, :3]).max()
    assert vrel <= 1e-4, vrel


def test_nbody3d_work_queue_is_split_independent_and_fits_small_scratch(gpus):
    """Partials are added in chunk order, and the chunking depends on n and the work-group size only: the same bits come out
    whether one device or all of them share the bodies.  A scratch array too small for that chunking widens the chunks
    (different rounding, same tolerance); one that cannot hold a single partial set per group makes the kernel refuse."""
    n = 16384
    pos0 = nbody3d_inputs(n)
    one, vone = run_nbody3dw(all_gpus()[0], pos0, 3, 4, 128)
    many, vmany = run_nbody3dw(all_gpus(), pos0, 3, 4, 128)
    assert np.array_equal(one, many) and np.array_equal(vone, vmany)
    again, _ = run_nbody3dw(all_gpus(), pos0, 3, 4, 128)
    assert np.array_equal(many, again)
    small, _ = run_nbody3dw(all_gpus()[0], pos0, 3, 4, 128, scratch_kfloats=(136 * 3 * 384 + 1023) // 1024 + 1)   # room for 3 chunks, not 8
    rel = np.abs(small[:, :3] - one[:, :3]) / np.maximum(np.abs(one[:, :3]), 1e-3)
    assert 0 < rel.max() <= 1e-4
    refused, _ = run_nbody3dw(all_gpus()[0], pos0, 1, 4, 128, scratch_kfloats=8)
    assert not refused.any()                                                    # untouched output, message on the device's stdout


def test_nbody3d_resident_steps_with_peer_stores(gpus):
    """nBody3Dwp: the step fused with the exchange of its results.  Positions are uploaded once; after that every device
    writes its share of new positions and velocities into ALL devices' buffers (NVLink peer stores), so the following steps
    move nothing over the host link, and only the last one downloads.  Must match the oracle like the host-mediated path,
    on one device (no peers to write to) as on several, and while the balancer moves the split."""
    n, steps, wpb, local = 16384, 6, 4, 128
    pos0 = nbody3d_inputs(n)
    params_h = np.array([1e-3, 1e-4, n, 0], np.float32)
    p, v, o = pos0.copy().reshape(-1), np.zeros(4 * n, np.float32), np.zeros(4 * n, np.float32)
    for _ in range(steps):
        po.nbody3d_step(p, v, o, params_h)
        p, o = o, p
    cr = ck.ClNumberCruncher(all_gpus(), kernels.source("nbody"))
    assert cr.errorCode() == 0, cr.errorMessage()
    kfloats = ((n // 128 + 8) * ((n // 256 + 7) // 8) * 384 + 4096) // 1024
    pa, pb, vel, params, scratch = (ck.ClFloatArray(4 * n), ck.ClFloatArray(4 * n), ck.ClFloatArray(4 * n), ck.ClFloatArray(8),
                                    ck.ClFloatArray(kfloats * 1024))
    pa.view[:] = pos0.reshape(-1)
    params.view[:5] = [1e-3, 1e-4, n, wpb, kfloats]
    params.readOnly = True
    scratch.read, scratch.write = False, False
    for a in (pa, pb, vel):
        a.numberOfElementsPerWorkItem = 4 // wpb
    tables = {id(pa): cr.peerTable(pa, vel), id(pb): cr.peerTable(pb, vel)}
    assert tables[id(pa)].view[0] == gpus and len(set(tables[id(pa)].view[1:1 + gpus].tolist())) == gpus
    src, dst = pa, pb
    for s in range(steps):
        last = s == steps - 1
        src.read, src.partialRead, src.write = (s == 0), False, False        # the only upload of positions
        dst.read, dst.partialRead, dst.write = False, False, last            # the only download
        vel.read, vel.partialRead, vel.write = False, False, last            # starts at zero in the (zero-filled) buffers
        host_before = dst.view.copy()
        assert src.nextParam(vel, dst, params, scratch, tables[id(dst)]).compute(cr, 1, "nBody3Dwp", n * wpb, local) == 0
        if not last:
            assert np.array_equal(dst.view, host_before) and not vel.view.any()   # nothing came back to the host in between
        src, dst = dst, src
    got, exp = src.view.reshape(n, 4), p.reshape(n, 4)
    rel = np.abs(got[:, :3] - exp[:, :3]) / np.maximum(np.abs(exp[:, :3]), 1e-3)
    assert rel.max() <= 1e-4, rel.max()
    assert np.array_equal(got[:, 3], exp[:, 3])
    vexp = v.reshape(n, 4)[:, :3]
    assert np.abs(vel.view.reshape(n, 4)[:, :3] - vexp).max() / np.abs(vexp).max() <= 1e-4


def oracle_sample_check(pos_before, vel_before, pos_after, vel_after, n, blocks=8, block=128):
    """One oracle step from (pos_before, vel_before) on `blocks` runs of `block` bodies spread over the whole range, against all
    n bodies, compared with the state the GPU produced: positions <= 1e-4 relative (BASELINE.json), masses carried over."""
    params_h = np.array([1e-3, 1e-4, n, 0], np.float32)
    checked = 0
    for b in range(blocks):
        first = min(n - block, (b * (n - block)) // max(1, blocks - 1) + (37 if b % 2 else 0))   # odd starts: not aligned to body groups
        v = vel_before.copy().reshape(-1)
        o = np.zeros(4 * n, np.float32)
        po.nbody3d_step(np.ascontiguousarray(pos_before.reshape(-1)), v, o, params_h, first, block)
        exp = o.reshape(n, 4)[first:first + block]
        got = pos_after[first:first + block]
        rel = np.abs(got[:, :3] - exp[:, :3]) / np.maximum(np.abs(exp[:, :3]), 1e-3)
        assert rel.max() <= 1e-4, (first, rel.max())
        assert np.array_equal(got[:, 3], exp[:, 3])
        vexp = v.reshape(n, 4)[first:first + block, :3]
        vgot = vel_after[first:first + block, :3]
        assert np.abs(vgot - vexp).max() <= 1e-4 * max(np.abs(vexp).max(), 1e-6), first
        checked += block
    return checked


@pytest.mark.parametrize("kernel", ["nBody3Dw", "nBody3Dwp"])
@pytest.mark.parametrize("devices", ["one", "all"])
def test_nbody3d_at_the_c4_size_against_the_oracle(gpus, devices, kernel):
    """BASELINE.json configs[3] at its stated size: 131,072 bodies, four work-items per body, work-groups of 128 (64 chunks of
    partial sums per body group, the 100 MB scratch).  The first and the tenth step are each checked against one oracle step
    taken from the state the GPU itself held before that step, on 1,024 bodies spread over the range (1.3e8 interactions on
    the host), so every step's kernel is verified without ten full host steps."""
    n, wpb, local, steps = 131072, 4, 128, 10
    devs = all_gpus()[0] if devices == "one" else all_gpus()
    pos0 = nbody3d_inputs(n, seed=0xC0FFEE)
    cr = ck.ClNumberCruncher(devs, kernels.source("nbody"))
    assert cr.errorCode() == 0, cr.errorMessage()
    groups, chunks = n // 128 + 8, (n // 256 + 7) // 8
    kfloats = (groups * chunks * 384 + groups + 256 + 1023) // 1024
    pa, pb, vel, params, scratch = (ck.ClFloatArray(4 * n), ck.ClFloatArray(4 * n), ck.ClFloatArray(4 * n), ck.ClFloatArray(8),
                                    ck.ClFloatArray(kfloats * 1024))
    pa.view[:] = pos0.reshape(-1)
    params.view[:5] = [1e-3, 1e-4, n, wpb, kfloats]
    params.readOnly = True
    scratch.read, scratch.write = False, False
    for a in (pa, pb, vel):
        a.numberOfElementsPerWorkItem = 4 // wpb
    fused = kernel == "nBody3Dwp"
    tables = {id(pa): cr.peerTable(pa, vel), id(pb): cr.peerTable(pb, vel)} if fused else None
    src, dst = pa, pb
    checked = 0
    for s in range(steps):
        look = s in (0, steps - 2, steps - 1)                                # states needed on the host: after steps 1, 9 and 10
        if fused:
            src.read, src.partialRead, src.write = (s == 0), False, False
            dst.read, dst.partialRead, dst.write = False, False, look
            vel.read, vel.partialRead, vel.write = False, False, look
            group = src.nextParam(vel, dst, params, scratch, tables[id(dst)])
        else:
            src.read, src.partialRead, src.write = True, False, False
            dst.read, dst.partialRead, dst.write = False, False, True
            vel.read, vel.partialRead, vel.write = False, True, True
            group = src.nextParam(vel, dst, params, scratch)
        before = (src.view.reshape(n, 4).copy(), vel.view.reshape(n, 4).copy()) if s in (0, steps - 1) else None
        assert group.compute(cr, 1, kernel, n * wpb, local) == 0
        if before is not None:
            checked += oracle_sample_check(before[0], before[1], dst.view.reshape(n, 4), vel.view.reshape(n, 4), n)
        src, dst = dst, src
    assert checked == 2 * 1024
    for a in (pa, pb, vel, params, scratch):
        a.dispose()
    cr.dispose()


# ------------------------------------------------------------------ blur → Sobel → histogram (K5)
@pytest.mark.parametrize("W,H", [(256, 128), (3840, 2160)])
def test_image_chain_bit_exact(gpus, W, H):
    n = W * H
    img = po.uniform(0xF00D, n, 0, 1)
    dims_h = np.array([W, H], np.int32)
    b_exp, s_exp, h_exp = np.zeros(n, np.float32), np.zeros(n, np.float32), np.zeros(256, np.uint32)
    po.blur5x5(img, b_exp, dims_h)
    po.sobel(b_exp, s_exp, dims_h)
    po.histogram(s_exp, h_exp)
    if (W, H) == (256, 128):
        g = GOLDEN["image_256x128"]
        assert sha(b_exp) == g["sha256_blur"] and sha(s_exp) == g["sha256_sobel"] and h_exp.tolist() == g["hist"]

    cr = ck.ClNumberCruncher(gpu0(), kernels.source("image"))
    assert cr.errorCode() == 0, cr.errorMessage()
    src, blur, sob, hist, dims = ck.ClFloatArray(n), ck.ClFloatArray(n), ck.ClFloatArray(n), ck.ClUIntArray(256), ck.ClIntArray(2)
    src.view[:] = img
    dims.view[:] = dims_h
    dims.readOnly = True
    src.readOnly = True                    # whole-array read: a stencil needs its neighbours' rows
    blur.writeOnly = True
    assert src.nextParam(blur, dims).compute(cr, 1, "gaussianBlur", n, 256) == 0
    assert np.array_equal(blur.view, b_exp)
    blur.writeOnly = False
    blur.readOnly = True
    sob.writeOnly = True
    assert blur.nextParam(sob, dims).compute(cr, 2, "sobel", n, 256) == 0
    assert np.array_equal(sob.view, s_exp)
    sob.writeOnly = False
    sob.readOnly = True
    sob.numberOfElementsPerWorkItem = 16
    hist.writeOnly = True
    hist.writeAll = True                   # the whole 256-bin array comes back from one device
    hist.view[:] = 12345
    for _ in range(2):                     # twice: the zeroing kernel makes the counts replace, not accumulate
        assert sob.nextParam(hist).compute(cr, 3, "histogramZero histogram", n // 16, 256) == 0
    assert np.array_equal(hist.view, h_exp)
    assert int(hist.view.sum()) == n       # a checksum that holds at any size


@pytest.mark.parametrize("W,H", [(256, 128), (3840, 2160), (24, 7)])
def test_gaussian_blur8_is_bit_identical_to_the_one_pixel_kernel(gpus, W, H):
    """gaussianBlur8: eight adjacent pixels per work-item with 16-byte loads; same operations per pixel, so the same bits
    as gaussianBlur and the oracle — split over every device (8 floats of `out` per work-item), including a frame so
    narrow that most work-items take the clamped edge path."""
    n = W * H
    img = po.uniform(0xBEEF, n, 0, 1)
    exp = np.zeros(n, np.float32)
    po.blur5x5(img, exp, np.array([W, H], np.int32))
    devs = all_gpus() if (n // 8) % (64 * gpus) == 0 else gpu0()
    cr = ck.ClNumberCruncher(devs, kernels.source("image"))
    assert cr.errorCode() == 0, cr.errorMessage()
    src, out, dims = ck.ClFloatArray(n), ck.ClFloatArray(n), ck.ClIntArray(2)
    src.view[:] = img
    dims.view[:] = [W, H]
    dims.readOnly = True
    src.readOnly = True
    out.writeOnly = True
    out.numberOfElementsPerWorkItem = 8
    local = 64 if (n // 8) % 64 == 0 else (n // 8)
    assert src.nextParam(out, dims).compute(cr, 1, "gaussianBlur8", n // 8, local) == 0
    assert np.array_equal(out.view, exp)


# ------------------------------------------------------------------ K5 strip kernels (one pass over the frame)
def image_oracle(img, W, H):
    n = W * H
    dims_h = np.array([W, H], np.int32)
    b, s, h = np.zeros(n, np.float32), np.zeros(n, np.float32), np.zeros(256, np.uint32)
    po.blur5x5(img, b, dims_h)
    po.sobel(b, s, dims_h)
    po.histogram(s, h)
    return b, s, h


def same_bits(a, b):
    return np.array_equal(a.view(np.uint32), b.view(np.uint32))


@pytest.mark.parametrize("W,H,rows", [(3840, 2160, 30), (3840, 2160, 59), (24, 7, 30), (24, 7, 2), (256, 128, 16), (1000, 333, 29),
                                      (124, 11, 3), (8, 5, 30), (4, 1, 1), (4096, 100, 30), (1920, 1080, 27)])
def test_strip_kernels_bit_exact(gpus, W, H, rows):
    """blurStrips, sobelStrips, histogramStrips and the fused blurSobelHistogram against oracle_blur5x5 / oracle_sobel /
    oracle_histogram: every pixel and every bin, frames with one strip and with many, bands that end inside the image, a
    band count that does not divide the rows, widths that end inside a strip.  Two launches in a row prove that the fused
    kernel's accumulator re-arms itself and that `hist` is overwritten, not added to."""
    n = W * H
    img = po.uniform(0xF00D + W, n, 0, 1)
    b_exp, s_exp, h_exp = image_oracle(img, W, H)
    cr = ck.ClNumberCruncher(gpu0(), kernels.source("image"))
    assert cr.errorCode() == 0, cr.errorMessage()
    src, blur, sob = ck.ClFloatArray(n), ck.ClFloatArray(n), ck.ClFloatArray(n)
    hist, acc, dims = ck.ClUIntArray(256), ck.ClUIntArray(2049), ck.ClIntArray(3)
    src.view[:] = img
    dims.view[:] = [W, H, rows]
    dims.readOnly = True
    src.readOnly = True
    acc.read, acc.write = False, False             # scratch accumulator: zero when the buffer is created, re-armed by the kernel
    g, l = kernels.strip_launch(W, H, rows)
    blur.view[:] = -1
    blur.read, blur.writeAll = False, True
    assert src.nextParam(blur, dims).compute(cr, 1, "blurStrips", g, l) == 0
    assert same_bits(blur.view, b_exp)
    blur.read, blur.write, blur.writeAll = True, False, False
    sob.view[:] = -1
    sob.read, sob.writeAll = False, True
    assert blur.nextParam(sob, dims).compute(cr, 2, "sobelStrips", g, l) == 0
    assert same_bits(sob.view, s_exp)
    hist.read, hist.writeAll = False, True
    for _ in range(2):
        hist.view[:] = 0xFFFFFFFF
        assert src.nextParam(hist, dims, acc).compute(cr, 3, "blurSobelHistogram", g, l) == 0
        assert np.array_equal(hist.view, h_exp)
    sob.read, sob.write, sob.writeAll = True, False, False
    for groups in (1, 37, 592):
        hist.view[:] = 0xFFFFFFFF
        assert sob.nextParam(hist, dims, acc).compute(cr, 4, "histogramStrips", groups * 256, 256) == 0
        assert np.array_equal(hist.view, h_exp)
    for a in (src, blur, sob, hist, acc, dims):
        a.dispose()
    cr.dispose()


def test_fused_strip_kernel_bins_special_values_like_the_oracle(gpus):
    """NaN, infinities, values whose gradients overflow, denormals and an all-constant frame (gradient exactly 0): the
    fused kernel's clamp-then-rsqrt binning must land every pixel in the oracle's bin."""
    W, H, rows = 3840, 270, 30
    n = W * H
    img = po.uniform(0xBEEF, n, 0, 1)
    img[5], img[n // 2], img[n // 3], img[n // 3 + 1], img[7 * n // 8] = np.nan, np.inf, 3e38, -3e38, 1e-30
    img[W * 100:W * 140] = 0.25                     # flat rows: m == 0 exactly
    img[W * 200 + 1000:W * 200 + 1400] = 1e-22      # gradients in the denormal range after squaring
    img[W * 220:W * 221] *= 40.0                    # magnitudes above 1: bin 255
    for frame in (img, np.zeros(n, np.float32), np.full(n, -2.5, np.float32)):
        _, _, h_exp = image_oracle(frame, W, H)
        cr = ck.ClNumberCruncher(gpu0(), kernels.source("image"))
        src, hist, acc, dims = ck.ClFloatArray(n), ck.ClUIntArray(256), ck.ClUIntArray(2049), ck.ClIntArray(3)
        src.view[:] = frame
        dims.view[:] = [W, H, rows]
        dims.readOnly = True
        src.readOnly = True
        acc.read, acc.write = False, False
        hist.read, hist.writeAll = False, True
        g, l = kernels.strip_launch(W, H, rows)
        assert src.nextParam(hist, dims, acc).compute(cr, 1, "blurSobelHistogram", g, l) == 0
        assert np.array_equal(hist.view, h_exp)
        assert int(hist.view.sum()) == n
        for a in (src, hist, acc, dims):
            a.dispose()
        cr.dispose()
