"""The reference's copy-kernel matrix (Tester.cs:32-6755, driver testTypesWithFeatures :6758-7067) replayed through the
C ABI: {C# array, C++ array} x {byte, char, int, uint, long, double, float} x {single, multi GPU} x
{no pipeline, event pipeline, driver pipeline} x {"test", "test test", "test test test"}.
The kernel text is the reference's own OpenCL C (`data2[id]=data[id]`), N = 1024, local 256 (plain) or 64 (pipelined),
4 blobs; the expected output is the input, bit-exact for every type (the reference allows ±1e-5 for floats)."""
import numpy as np
import pytest

import cekirdekler_b200 as ck
import reference_kernels
from cekirdekler_b200 import kernels
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu

N = 1024
TYPES = {
    # name: (OpenCL C element type in the reference kernel, numpy dtype, fill value used by Tester.cs)
    "Byte": ("uchar", np.uint8, 55),
    "Char": ("char16", np.uint16, ord("a")),   # C# char = UTF-16 code unit: the reference copies it as 2 uchars (Tester.cs:117-121)
    "Int": ("int", np.int32, -55),
    "UInt": ("uint", np.uint32, 55),
    "Long": ("long", np.int64, -55),
    "Double": ("double", np.float64, 0.001),
    "Float": ("float", np.float32, 0.001),
}
_crunchers = {}


def cruncher(tname, multi, gpus):
    key = (tname, multi)
    if key not in _crunchers:
        devs = ck.ClPlatforms.all().gpus()
        devs = devs if multi else devs[0]
        src = reference_kernels.REFERENCE_COPY_KERNEL_CHAR if tname == "Char" else reference_kernels.reference_copy_kernel(TYPES[tname][0])
        cr = ck.ClNumberCruncher(devs, src)
        assert cr.errorCode() == 0, cr.errorMessage()
        _crunchers[key] = cr
    return _crunchers[key]


def make_array(kind, dtype, fill):
    if kind == "fast":
        a = ck.ClArray(N, dtype)           # C++ array (FastArr): pinned native memory
        a.view[:] = fill
        return a
    host = np.full(N, fill, dtype)          # C# array: plain language array
    return ck.ClArray(host)


@pytest.mark.parametrize("flags", ["defaults", "partial"])
@pytest.mark.parametrize("kernel_list", ["test", "test test", "test test test"])
@pytest.mark.parametrize("pipe", ["none", "event", "driver"])
@pytest.mark.parametrize("multi", [False, True], ids=["SingleGPU", "MultiGPU"])
@pytest.mark.parametrize("tname", list(TYPES))
@pytest.mark.parametrize("kind", ["simple", "fast"])
def test_copy_matrix(gpus, kind, tname, multi, pipe, kernel_list, flags):
    ctype, dtype, fill = TYPES[tname]
    cr = cruncher(tname, multi, gpus)
    data = make_array(kind, dtype, fill)
    # make the input position-dependent as well, so offset/range mistakes cannot hide behind a constant fill
    data.view[:] = (np.arange(N) % 97).astype(dtype) + np.dtype(dtype).type(fill)
    data2 = make_array(kind, dtype, 0)
    data2.view[:] = 0
    if flags == "partial":
        # sliced uploads (pipelinable); "defaults" keeps the reference tests' flags: whole read + sliced write
        data.partialRead = True
        data.read = False
        data2.partialRead = True
        data2.read = False
    cid = {"none": 1, "event": 2, "driver": 3}[pipe] * 10 + len(kernel_list.split())
    g = data.nextParam(data2)
    if pipe == "none":
        rc = g.compute(cr, cid, kernel_list, N, 256)
    else:
        rc = g.compute(cr, cid, kernel_list, N, 64, 0, True, pipe == "event", 4)
    needed = cr.numberOfDevices * (256 if pipe == "none" else 64 * 4)
    if multi and needed > N:
        assert rc != 0  # the reference refuses: global range < devices*local(*blobs) (ClArray.cs:1640-1657)
        cr_err = cr.numberOfErrorsHappened
        assert cr_err >= 1
        _crunchers.pop((tname, multi))  # a cruncher that has errored stays inert, as in the reference
        return
    assert rc == 0
    expect = np.zeros(N, dtype)
    po.copy(data.view, expect)
    assert data2.view.tobytes() == expect.tobytes()
    assert data2.view.tobytes() == data.view.tobytes()


def test_copy_matrix_runs_twice_with_rebalanced_ranges(gpus):
    """Repeated computes on one compute id re-balance the split; results must stay exact (Tester runs each case once,
    the nBody test 150 times)."""
    cr = cruncher("Int", True, gpus)
    a, b = ck.ClArray(1 << 16, np.int32), ck.ClArray(1 << 16, np.int32)
    a.view[:] = np.arange(1 << 16, dtype=np.int32) * 7 - 3
    a.partialRead, a.read = True, False
    for _ in range(12):
        b.view[:] = 0
        assert a.nextParam(b).compute(cr, 99, "test", 1 << 16, 256) == 0
        assert (b.view == a.view).all()
        r = cr.globalRanges(99)
        assert sum(r) == 1 << 16 and all(x % 256 == 0 for x in r)
