"""Kernel-source front-end on a CPU-only box: the signature rewrite, and that NVRTC builds every shipped kernel source
(and the reference's own kernel strings) for sm_100a without a device."""
import ctypes as C

import pytest

from cekirdekler_b200 import kernels
import reference_kernels
from cekirdekler_b200._native import shim


def rewrite(src: str) -> str:
    L = shim()
    L.ckRewriteKernelSource.restype = C.c_void_p
    L.ckRewriteKernelSource.argtypes = [C.c_char_p]
    p = L.ckRewriteKernelSource(src.encode())
    try:
        return C.string_at(p).decode()
    finally:
        C.CDLL(None).free(C.c_void_p(p))


def compile_check(src: str):
    L = shim()
    L.ckCompileCheck.argtypes = [C.c_char_p, C.c_int, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_longlong)]
    log, nbytes = C.c_char_p(), C.c_longlong()
    rc = L.ckCompileCheck(src.encode(), 10, 0, C.byref(log), C.byref(nbytes))
    return rc, (log.value or b"").decode(errors="replace"), nbytes.value


def test_rewrite_adds_the_hidden_offset_parameter():
    out = rewrite("__kernel void f(__global float* a, __global int* b) { a[get_global_id(0)] = b[0]; }")
    assert 'extern "C" __global__ void f(const unsigned int __ck_goff, __global float* a, __global int* b)' in out
    out = rewrite("kernel void g(void) { }\nkernel void h() { }")
    assert "void g(const unsigned int __ck_goff)" in out and "void h(const unsigned int __ck_goff)" in out


def test_rewrite_leaves_comments_strings_and_lookalikes_alone():
    src = '// __kernel void c(int x)\n/* kernel void d(int x) */\nconst char* s = "kernel void e(";\nint my_kernel = 3; void kernelize(int q);'
    assert rewrite(src) == src


def test_rewrite_turns_vector_literals_into_make_calls():
    assert rewrite("float4 v = (float4)(a, b, c, d);") == "float4 v = make_float4(a, b, c, d);"
    assert rewrite("x = ( uint2 ) (1, 2) + (int3)(0);") == "x = make_uint2(1, 2) + make_int3(0);"
    # casts, lookalikes, comments and strings stay as they are
    for src in ["y = (float4)v;", "y = (float)(a + b);", "y = (float5)(1);", "y = (myfloat4)(1);", "// (float4)(1,2,3,4)",
                'printf("(float4)(%f)", q);']:
        assert rewrite(src) == src


def test_vector_literals_and_more_built_ins_compile():
    src = """
    float3 reflect3(float3 i, float3 n) { return i - 2.0f * dot(n, i) * n; }
    __kernel void k(__global const float* in, __global float* out, __global uint* bits)
    {
        size_t i = get_global_id(0);
        float4 v = vload4(i, in);
        float4 w = (float4)(v.x, (float2)(1.0f, 2.0f), 0.5f) * (float4)(2.0f) + (float4)((float3)(v.y), v.w);
        w = clamp(mix(w, v, 0.25f), 0.0f, 10.0f) / (1.0f + fabs(w));
        w -= normalize(v) * length(w.x * v);
        float3 r = reflect3((float3)(w.x, w.y, w.z), normalize(cross((float3)(1.0f, 0.0f, 0.0f), (float3)(v.x, v.y, 1.0f))));
        int4 q = convert_int4(w) + (int4)(1, 2, 3, 4);
        vstore4(mad(w, v, (float4)(r, distance(w, v))) + convert_float4(q), i, out);
        bits[i] = rotate((uint)q.x, 3u) ^ popcount((uint)q.y) ^ mul24(q.z, q.w) ^ (uint)select(1, 2, q.x > 0);
        out[4 * i] += smoothstep(0.0f, 1.0f, fract(v.x)) + step(0.5f, v.y) + radians(degrees(v.z));
    }"""
    rc, log, _ = compile_check(src)
    assert rc == 0, log


@pytest.mark.parametrize("name", ["stream", "mandelbrot", "nbody", "image"])
def test_shipped_kernels_build_for_sm_100a(name):
    rc, log, nbytes = compile_check(kernels.source(name))
    assert rc == 0, log
    assert nbytes > 1000


@pytest.mark.parametrize("src", [
    reference_kernels.reference_copy_kernel("uchar"), reference_kernels.reference_copy_kernel("double"), reference_kernels.reference_copy_kernel("long"),
    reference_kernels.REFERENCE_COPY_KERNEL_CHAR, reference_kernels.REFERENCE_VECTOR_ADD, reference_kernels.reference_nbody_kernel(8192),
    reference_kernels.REFERENCE_WAVE_EQUATION,
])
def test_reference_opencl_strings_build_unmodified(src):
    rc, log, _ = compile_check(src)
    assert rc == 0, log


def test_opencl_c_idioms_compile_through_the_prelude():
    src = """
    float helper(float x) { return mad(x, 2.0f, 1.0f); }
    __kernel void k(__global float* a, __global const float* b, __global int* hist)
    {
        __local float tile[64];
        int l = get_local_id(0);
        int i = get_global_id(0);
        tile[l] = b[i];
        barrier(CLK_LOCAL_MEM_FENCE);
        float4 v = make_float4(tile[l], 1.0f, 2.0f, 3.0f);
        v = v * 2.0f + v;
        a[i] = helper(native_sqrt(dot(v, v))) + (float)get_group_id(0) + (float)get_local_size(0);
        atomic_add(&hist[i % 16], 1);
        atomic_inc(&hist[0]);
    }"""
    rc, log, _ = compile_check(src)
    assert rc == 0, log


def test_build_errors_are_reported_not_thrown():
    rc, log, nbytes = compile_check("__kernel void broken(__global float* a) { a[0] = undefined_symbol; }")
    assert rc != 0 and "undefined_symbol" in log and nbytes == 0


def test_component_selections_and_the_readme_sample_compile():
    """`.xy .zw .lo .hi .xyz .s0-.s3` on values and as assignment targets (OpenCL C 6.1.7), a user struct that has a member
    called `lo` of its own, and the README's first example (/root/reference/README.md:52-59)."""
    src = """
    struct Range { float lo; float hi; };
    __kernel void swz(__global float4* v, __global float2* o, __global float* s)
    {
        int i = get_global_id(0);
        float4 a = v[i];
        float2 p = a.xy + a.zw;
        float2 q = v[i].lo * 2.0f - (a * 2.0f).hi;
        a.zw = p;
        a.lo = (float2)(a.s3, a.s0);
        float3 t = a.xyz;
        struct Range r; r.lo = t.x; r.hi = t.y + 1.5f;
        o[i] = p + q + (float2)(r.lo, r.hi);
        s[i] = a.s0 + a.s1 + a.s2 + a.s3 + p.lo + q.hi + 1.5f;
        v[i] = a;
    }
    __kernel void hello(__global char * arr)
    {
        printf("hello world");
    }"""
    rc, log, _ = compile_check(src)
    assert rc == 0, log


def test_device_side_enqueue_is_refused_with_a_clear_message():
    src = """
    __kernel void child(__global int* a) { a[get_global_id(0)] = 1; }
    __kernel void parent(__global int* a)
    {
        enqueue_kernel(get_default_queue(), CLK_ENQUEUE_FLAGS_NO_WAIT, ndrange_1D(64), ^{ child(a); });
    }"""
    rc, log, nbytes = compile_check(src)
    assert rc != 0 and nbytes == 0
    assert "enqueue_kernel" in log and "not supported" in log


def test_component_selection_rewrite_on_postfix_expressions():
    """What the source rewrite does with `.xy .zw .lo .hi .xyz .s0-.s3`: the whole postfix expression in front is wrapped,
    numbers and comments are left alone."""
    from cekirdekler_b200 import kernels

    def body(src):
        return kernels.rewritten_with_prelude(src).split('#line 1 "user_kernels.cu"\n', 1)[1]

    out = body("void f(float4 a, float4* p, S* q) { float2 r = a.xy; r = p[3].zw; r = q->v.lo; r = g(a, 1).hi; float3 t = (a + a).xyz; "
               "float s = a.s2 + 1.5f + 2.f; /* a.xy stays */ }")
    assert "ck_xy(a)" in out and "ck_zw(p[3])" in out and "ck_lo(q->v)" in out and "ck_hi(g(a, 1))" in out and "ck_xyz((a + a))" in out
    assert "a.z + 1.5f + 2.f" in out and "/* a.xy stays */" in out
    assert body("int hi = lo.hi;") == "int hi = ck_hi(lo);"           # names called lo / hi are ordinary identifiers
    assert body("x = v2.xy;") == "x = ck_xy(v2);"                     # an identifier that ends in a digit is not a number
