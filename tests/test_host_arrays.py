"""Host-array behaviour the reference's Tester pins with its *ArrayOperations() methods (Tester.cs:7076-7672):
indexing, conversion from language arrays, CopyFrom/CopyTo, and the flag rules of ClArray (ClArray.cs:837-845, :1742-1863)."""
import numpy as np
import pytest

import cekirdekler_b200 as ck

TYPES = [
    (ck.ClByteArray, np.uint8, 55), (ck.ClCharArray, np.uint16, ord("a")), (ck.ClIntArray, np.int32, -55),
    (ck.ClUIntArray, np.uint32, 55), (ck.ClLongArray, np.int64, -55), (ck.ClDoubleArray, np.float64, 0.001),
    (ck.ClFloatArray, np.float32, 0.001),
]


@pytest.mark.parametrize("ctor,dtype,fill", TYPES)
def test_fast_array_operations(ctor, dtype, fill):
    a = ctor(1000)
    assert a.Length == 1000 and a.fastArr and a.dtype == np.dtype(dtype)
    assert a.view.ctypes.data % 4096 == 0  # default alignment (ClArray.cs:845)
    assert (a.ToArray() == 0).all()
    for i in range(1000):
        a[i] = fill
    assert a[500] == np.dtype(dtype).type(fill)
    b = np.zeros(1000, dtype)
    a.CopyTo(b)
    assert (b == a.view).all()
    c = ctor(1000)
    c.CopyFrom(b)
    assert (c.view == b).all()
    a.dispose()
    a.dispose()  # callable multiple times (ClArray.cs:848-850)


@pytest.mark.parametrize("ctor,dtype,fill", TYPES)
def test_switching_representation_and_resizing_keep_values_and_flags(ctor, dtype, fill):
    """The tail of the reference's *ArrayOperations tests (Tester.cs:7465-7483): `fastArr = false`, then CopyFrom/CopyTo
    must still round-trip; plus the N setter (ClArray.cs:749-800): shorter keeps a prefix, longer pads with zeros."""
    a = ctor(1024)
    a.view[:] = (np.arange(1024) % 100).astype(dtype)
    a.partialRead, a.read, a.numberOfElementsPerWorkItem = True, False, 4
    before = a.view.copy()
    a.fastArr = False
    assert not a.fastArr and a.Length == 1024 and (a.view == before).all()
    assert a.partialRead and not a.read and a.write and a.numberOfElementsPerWorkItem == 4
    src = ((35 * np.arange(1024)) % 251).astype(dtype)
    dst = np.zeros(1024, dtype)
    a.CopyFrom(src)
    a.CopyTo(dst)
    assert (dst == src).all()
    a.fastArr = True
    assert a.fastArr and (a.view == src).all() and a.view.ctypes.data % 4096 == 0
    a.N = 300
    assert a.N == 300 and a.Length == 300 and a.fastArr and (a.view == src[:300]).all()
    a.N = 2000
    assert a.N == 2000 and (a.view[:300] == src[:300]).all() and not a.view[300:].any()
    assert a.partialRead and a.numberOfElementsPerWorkItem == 4
    ro = ctor(64)
    ro.readOnly = True
    ro.fastArr = False
    assert ro.readOnly and not ro.write and not ro.fastArr
    a.dispose()
    ro.dispose()


def test_alignment_parameter():
    a = ck.ClArray(100, np.float32, alignmentBytes=65536)
    assert a.view.ctypes.data % 65536 == 0


def test_wrap_language_array_shares_memory():
    x = np.arange(16, dtype=np.float32)
    a = ck.ClArray(x)
    assert not a.fastArr and a.Length == 16
    a[3] = 99
    assert x[3] == 99


def test_flag_defaults_and_exclusion_rules():
    a = ck.ClArray(8, np.float32)
    # defaults (ClArray.cs:837-845)
    assert (a.read, a.partialRead, a.write, a.writeAll, a.readOnly, a.writeOnly, a.zeroCopy) == (True, False, True, False, False, False, False)
    assert a.numberOfElementsPerWorkItem == 1 and a.alignmentBytes == 4096
    assert a.flagString() == " read  write "
    a.readOnly = True  # clears the write flags (ClArray.cs:1750-1765)
    assert (a.write, a.writeAll, a.readOnly) == (False, False, True)
    a.write = True     # refused while readOnly
    assert not a.write
    a.writeOnly = True  # refused while readOnly (ClArray.cs:1774)
    assert not a.writeOnly
    a.readOnly = False
    a.writeOnly = True  # clears the read flags
    assert (a.read, a.partialRead, a.writeOnly) == (False, False, True)
    a.read = True
    assert not a.read
    a.write = True
    a.zeroCopy = True
    assert a.flagString() == " write  wo  zc "


def test_flag_string_order():
    a = ck.ClArray(8, np.int32)
    a.partialRead = True
    a.writeAll = True
    assert a.flagString() == " partial  read  write  all "


def test_parameter_group_snapshots_flags_and_raw_array_defaults():
    a = ck.ClArray(8, np.float32)
    raw1, raw2 = np.zeros(8, np.float32), np.zeros(8, np.float32)
    g = a.nextParam(raw1)          # ClArray.nextParam(Array[]): raw arrays default to partialRead = false
    assert g._flags[1].split() == ["read", "write"]
    g2 = g.nextParam(raw2)         # ClParameterGroup.nextParam(Array[]): partialRead = true (quirk Q11)
    assert g2._flags[2].split() == ["partial", "read", "write"]
    a.partialRead = True           # changing the array after chaining does not change the group (flags were copied)
    assert g._flags[0].split() == ["read", "write"]


def test_kernel_name_extraction_matches_reference_regex():
    src = """
        __kernel void hello(__global char * arr) { }
        kernel   void  second_one (global float* a){}
        void helper(float x) {}
    """
    assert ck.ClNumberCruncher.kernelNames(src) == ["hello", "second_one"]
    assert ck.ClNumberCruncher.kernelNames("float f(float x){return x;}") == []


def test_device_selection_algebra():
    d = ck.ClDevices([0, 1, 2, 3])
    assert d.Length == 4 and d[2].ordinals == [2]
    assert (d[0] + d[3]).ordinals == [0, 3]       # ClDevices operator+ (ClObjectApi.cs:813-829)
    assert d.cpus().Length == 0 and d.accelerators().Length == 0
    assert d.gpus(streamingEnabled=True).streamingEnabled
    assert ck.ClPlatforms.all().Length == 1


def test_cruncher_without_kernels_or_devices_is_inert():
    # no kernel definitions → errorCode 1, nothing throws (ClNumberCruncher.cs:323-338)
    cr = ck.ClNumberCruncher(ck.ClDevices([0]), "float f(float x){return x;}")
    assert cr.errorCode() == 1 and "no kernel definitions" in cr.errorMessage()
    a = ck.ClArray(1024, np.float32)
    assert a.compute(cr, 1, "f", 1024, 256) != 0   # refused, not raised
    assert cr.numberOfErrorsHappened >= 1


def test_numa_preference_helper_never_raises():
    from cekirdekler_b200 import spmd

    info = spmd.prefer_host_memory_near_gpu(0)       # no NVML / single node here: reports and leaves the policy alone
    assert set(info) >= {"numa_node", "applied"} and info["applied"] in (True, False)
