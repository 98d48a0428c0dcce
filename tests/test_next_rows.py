"""the Unity sample's waveEquation kernel on arrays of structs (Kamera.cs:234-268, ClArray.cs:1058-1074),
OpenCL-C idioms running unmodified through the NVRTC prelude, and the micro-benchmark device sorter (ClObjectApi.cs:1222-1244)."""
import numpy as np
import pytest

import cekirdekler_b200 as ck
import reference_kernels
from cekirdekler_b200 import kernels
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu


def test_wave_equation_on_wrapped_structs(gpus):
    """Kamera.cs: Vector3[] vertices/normals wrapped as byte arrays (12 bytes per work-item), `arguments` a plain float array;
    the kernel text is the reference's own; the oracle is the reference's CPU branch (Kamera.cs:213-218)."""
    n = 224 * 256
    rng = np.random.default_rng(7)
    base = rng.uniform(-1, 1, (n, 3)).astype(np.float32)
    normals = base / np.maximum(np.linalg.norm(base, axis=1, keepdims=True), 1e-6).astype(np.float32)
    verts = np.zeros((n, 3), np.float32)
    xyz, xyzn, xyzo = ck.ClArray.wrapArrayOfStructs(base), ck.ClArray.wrapArrayOfStructs(normals), ck.ClArray.wrapArrayOfStructs(verts)
    assert xyz.numberOfElementsPerWorkItem == 12 and xyz.Length == n * 12
    args_h = np.zeros(64, np.float32)
    arguments = ck.ClArray(args_h)
    arguments.write, arguments.partialRead = False, False          # Kamera.cs:255
    for a in (xyz, xyzn):
        a.partialRead, a.read, a.write = True, False, False
    xyzo.read = False
    cr = ck.ClNumberCruncher(ck.ClPlatforms.all().gpus(), reference_kernels.REFERENCE_WAVE_EQUATION)
    assert cr.errorCode() == 0, cr.errorMessage()
    for frame in range(4):
        args_h[:5] = [1.0 + 0.1 * frame, 0.05 * frame, base[0, 0], base[0, 1], n]
        assert xyz.nextParam(xyzn, xyzo, arguments).compute(cr, 1, "waveEquation", n, 64) == 0
        exp = np.zeros((n, 3), np.float32)
        po.wave_equation(base.reshape(-1), normals.reshape(-1), exp.reshape(-1), args_h)
        assert np.allclose(verts, exp, rtol=0, atol=2e-5), np.abs(verts - exp).max()


def test_opencl_c_idioms_run(gpus):
    src = """
    float helper(float x) { return mad(x, 2.0f, 1.0f); }
    __kernel void idioms(__global const float* in, __global float* out, __global int* hist, __global int* gid, __global int* ggid)
    {
        __local float tile[256];
        int l = get_local_id(0);
        int i = get_global_id(0);
        tile[l] = in[i];
        barrier(CLK_LOCAL_MEM_FENCE);
        float left = tile[(l + get_local_size(0) - 1) % get_local_size(0)];
        float4 v = (float4)((float2)(tile[l], left), 1.0f, 0.0f);
        v = v * (float4)(2.0f) + v;
        out[i] = helper(dot(v, v)) + (float)((get_global_id(0) / get_local_size(0)) % 2);
        gid[i] = (int)get_group_id(0);             // OpenCL: counted from the start of this enqueue, i.e. of the device's share
        ggid[i] = (int)ck_global_group_id();       // extension: counted from the start of the whole range
        atomic_add(&hist[i % 16], 1);
    }"""
    n = 1 << 14
    cr = ck.ClNumberCruncher(ck.ClPlatforms.all().gpus(), src)
    assert cr.errorCode() == 0, cr.errorMessage()
    a, out, hist = ck.ClArray(n, np.float32), ck.ClArray(n, np.float32), ck.ClArray(16, np.int32)
    gid, ggid = ck.ClArray(n, np.int32), ck.ClArray(n, np.int32)
    a.view[:] = np.arange(n) % 17
    a.readOnly = True
    out.writeOnly = True
    gid.writeOnly = True
    ggid.writeOnly = True
    hist.writeAll = True
    assert a.nextParam(out, hist, gid, ggid).compute(cr, 1, "idioms", n, 256) == 0
    assert np.array_equal(ggid.view, np.arange(n) // 256)
    for ref, rng in zip(cr.globalReferences(1), cr.globalRanges(1)):
        assert np.array_equal(gid.view[ref:ref + rng], np.arange(rng) // 256)
    x = a.view.reshape(-1, 256)
    left = np.roll(x, 1, axis=1)
    v2 = (3 * x) ** 2 + (3 * left) ** 2 + 9.0
    group = (np.arange(n) // 256 % 2).reshape(-1, 256)
    assert np.allclose(out.view.reshape(-1, 256), v2 * 2 + 1 + group, rtol=1e-6)
    if gpus == 1:
        assert hist.view.sum() == n and (hist.view == n // 16).all()


def test_devices_sorted_by_nbody_micro_benchmark(gpus):
    devs = ck.ClPlatforms.all().gpus()
    ranked = devs.devicesWithHighestDirectNbodyPerformance(4096)
    assert sorted(ranked.ordinals) == sorted(devs.ordinals)
    if gpus < 2:
        return
    # the order must follow the measurement: keep device 0 busy with someone else's kernels while the ranking runs,
    # and it has to come out last
    import threading

    load = ck.ClNumberCruncher(devs[0], """
        __kernel void burn(__global float* a) { int i = get_global_id(0); float v = a[i];
            for (int k = 0; k < 40000; k++) v = v * 1.0000001f + 1e-7f; a[i] = v; }""")
    assert load.errorCode() == 0, load.errorMessage()
    a = ck.ClArray(1 << 20, np.float32)
    a.read, a.write = False, False
    stop = threading.Event()

    def burn():
        while not stop.is_set():
            a.compute(load, 1, "burn", 1 << 20, 256)

    th = threading.Thread(target=burn)
    th.start()
    try:
        ranked = devs.devicesWithHighestDirectNbodyPerformance(16384)
    finally:
        stop.set()
        th.join()
    assert ranked.ordinals[-1] == devs.ordinals[0], ranked.ordinals
    load.dispose()


def test_compute_follows_an_array_that_was_resized_or_switched(gpus):
    """ClArray.N and ClArray.fastArr replace the backing store (ClArray.cs:749-800, :889-958); the next compute() must
    use the new one, on every device of the split."""
    cr = ck.ClNumberCruncher(ck.ClPlatforms.all().gpus(), reference_kernels.reference_copy_kernel("float"))
    assert cr.errorCode() == 0, cr.errorMessage()
    n = 256 * 64
    a, b = ck.ClFloatArray(n), ck.ClFloatArray(n)
    a.view[:] = np.arange(n)
    assert a.nextParam(b).compute(cr, 1, "test", n, 256) == 0
    assert np.array_equal(b.view, a.view)
    for arr in (a, b):
        arr.N = 2 * n
    a.view[n:] = -np.arange(n)
    a.fastArr = False
    assert a.nextParam(b).compute(cr, 1, "test", 2 * n, 256) == 0
    assert np.array_equal(b.view, a.view) and b.view[n + 5] == -5
    b.fastArr = False
    b.view[:] = 0
    assert a.nextParam(b).compute(cr, 1, "test", 2 * n, 256) == 0
    assert np.array_equal(b.view, a.view)


def test_component_selections_compute_what_opencl_says(gpus):
    """§8(f) row 1: `.xy .zw .lo .hi .xyz .s0-.s3` as values and as assignment targets."""
    src = """
    __kernel void swz(__global float4* v, __global float2* o, __global float* s)
    {
        int i = get_global_id(0);
        float4 a = v[i];
        float2 p = a.xy + a.zw;
        float2 q = v[i].lo * 2.0f - (a * 2.0f).hi;
        a.zw = p;
        a.lo = (float2)(a.s3, a.s0);
        float3 t = a.xyz;
        o[i] = p + q + (float2)(t.x, t.z);
        s[i] = a.s0 + a.s1 + a.s2 + a.s3 + p.lo + q.hi;
        v[i] = a;
    }"""
    n = 1 << 12
    cr = ck.ClNumberCruncher(ck.ClPlatforms.all().gpus(), src)
    assert cr.errorCode() == 0, cr.errorMessage()
    v, o, s = ck.ClArray(4 * n, np.float32), ck.ClArray(2 * n, np.float32), ck.ClArray(n, np.float32)
    v0 = (np.arange(4 * n, dtype=np.float32) % 13).reshape(n, 4)
    v.view[:] = v0.reshape(-1)
    v.numberOfElementsPerWorkItem, o.numberOfElementsPerWorkItem = 4, 2
    v.partialRead, v.read = True, False
    o.writeOnly, s.writeOnly = True, True
    assert v.nextParam(o, s).compute(cr, 1, "swz", n, 256) == 0
    x, y, z, w = v0[:, 0], v0[:, 1], v0[:, 2], v0[:, 3]
    p = np.stack([x + z, y + w], 1)
    q = np.stack([2 * x - 2 * z, 2 * y - 2 * w], 1)
    a = np.stack([p[:, 1], x, p[:, 0], p[:, 1]], 1)            # zw = p, then lo = (s3, s0) = (p.y, x)
    assert np.array_equal(v.view.reshape(n, 4), a)
    assert np.array_equal(o.view.reshape(n, 2), p + q + np.stack([a[:, 0], a[:, 2]], 1))
    assert np.array_equal(s.view, a.sum(1) + p[:, 0] + q[:, 1])


def test_readme_hello_world_prints_from_every_work_item(gpus, capfd):
    """The README's first example (/root/reference/README.md:52-59): 1000 work-items in groups of 100, shared by all GPUs,
    each printing "hello world" through the device's printf."""
    src = """
                    __kernel void hello(__global char * arr)
                    {
                        printf("hello world");
                    }
                """
    cr = ck.ClNumberCruncher(ck.ClPlatforms.all().gpus(), src)
    assert cr.errorCode() == 0, cr.errorMessage()
    array = ck.ClArray(1000, np.uint8)
    assert array.compute(cr, 1, "hello", 1000, 100) == 0
    cr.sync()
    cr.dispose()                         # the device's printf buffer is flushed at synchronisation points
    out = capfd.readouterr().out
    assert out.count("hello world") == 1000
