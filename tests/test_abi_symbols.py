"""The C-ABI libraries load on a CPU-only box and export every symbol their headers declare (no compute calls)."""
import ctypes
import re
import subprocess
from pathlib import Path

from cekirdekler_b200._native import abi_symbols, host, shim

ROOT = Path(__file__).resolve().parent.parent

# the 81 names the reference's C# binds, by group
REFERENCE_81 = """
platformList numberOfPlatforms deletePlatformList
createPlatform deletePlatform getPlatformInfo getPlatformNameString getPlatformVendorNameString CODE_CPU CODE_GPU CODE_ACC
numberOfCpusInPlatform numberOfGpusInPlatform numberOfAcceleratorsInPlatform
createDevice createDeviceAsPartition deleteDevice getDeviceName getDeviceVendorName deviceGDDR deviceComputeUnits deviceMemSize
createContext deleteContext
createCommandQueue createCommandQueue2 deleteCommandQueue addMarkerToCommandQueue getMarkerCounterOfCommandQueue resetMarkerCounterOfCommandQueue
createEvent deleteEvent
createEventArr addToEventArr deleteEventArr
createUserEvent deleteUserEvent triggerUserEvent addUserEvent decrementUserEvent incrementUserEvent
createString writeToString deleteString readFromString jsonStringCallBack
createProgram deleteProgram getProgramErr readProgramErrString
createKernel deleteKernel getKernelErr setKernelArgument setKernelArguments
createNdRange deleteNdRange
compute computeRepeated computeRepeatedWithSyncKernel computeEvent
finish flush wait2 wait3 waitN
createBuffer deleteBuffer
writeToBuffer writeToBufferEvent writeToBufferRanged writeToBufferRangedEvent
readFromBuffer readFromBufferEvent readFromBufferRanged readFromBufferRangedEvent
sizeOf createArray alignedArrHead deleteArray copyMemory
""".split()


def test_reference_abi_has_81_symbols():
    assert len(REFERENCE_81) == 81 and len(set(REFERENCE_81)) == 81


def test_shim_exports_every_declared_symbol():
    L = shim()
    declared = abi_symbols()
    assert set(REFERENCE_81) <= set(declared)
    missing = [n for n in declared if not hasattr(L, n)]
    assert not missing, missing


def test_host_exports_every_declared_symbol():
    L = host()
    hdr = (ROOT / "include" / "cekirdekler_host.h").read_text()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = set(re.findall(r"\b(ckh_[a-z0-9_]+)\s*\(", hdr))
    assert len(names) > 30
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing


def test_shim_has_no_hard_driver_dependency():
    # the library must load without libcuda.so.1 (the CPU build/test box has no driver): cudart is linked statically
    out = subprocess.run(["ldd", str(ROOT / "cekirdekler_b200" / "lib" / "libKutuphaneCL.so")], capture_output=True, text=True).stdout
    assert "libcuda.so" not in out


def test_codes_sizes_and_platform_without_gpu():
    L = shim()
    assert (L.CODE_CPU(), L.CODE_GPU(), L.CODE_ACC()) == (2, 4, 8)
    table = (ctypes.c_int32 * 20)()
    L.sizeOf(table)
    # CSpaceArrays.ARR_*: FLOAT, DOUBLE, LONG, INT, UINT, BYTE, CHAR (CSpaceArrays.cs:48-78)
    assert list(table[:7]) == [4, 8, 8, 4, 4, 1, 2]
    lst = L.platformList()
    assert L.numberOfPlatforms(lst) == 1
    plat = L.createPlatform(lst, 0)
    L.deletePlatformList(lst)
    assert L.readFromString(L.getPlatformNameString(plat)) == b"NVIDIA CUDA"
    assert L.numberOfCpusInPlatform(plat) == 0 and L.numberOfAcceleratorsInPlatform(plat) == 0
    assert L.numberOfGpusInPlatform(plat) == L.ckDeviceCount()
    L.deletePlatform(plat)
    L.deletePlatform(plat)  # a repeated delete is a no-op (finalizer-order safety)


def test_handles_survive_out_of_order_deletes():
    L = shim()
    s = L.createString()
    L.writeToString(s, b"hello")
    assert L.readFromString(s) == b"hello"
    L.jsonStringCallBack(s)  # releases
    L.deleteString(s)        # stale handle: ignored
    r = L.createNdRange(7)
    L.deleteNdRange(r)
    L.deleteNdRange(r)
    e = L.createEvent()
    arr = L.createEventArr(0)
    L.addToEventArr(arr, e, 0)
    L.deleteEvent(e)       # the array still holds a reference
    L.deleteEventArr(arr)
    L.deleteEventArr(arr)


def test_copy_memory_and_arrays_without_gpu():
    L = shim()
    a = L.createArray(100, 4096, 3)  # ARR_INT
    b = L.createArray(100, 4096, 3)
    pa, pb = L.alignedArrHead(a), L.alignedArrHead(b)
    assert pa % 4096 == 0 and pb % 4096 == 0
    ia = (ctypes.c_int32 * 100).from_address(pa)
    ib = (ctypes.c_int32 * 100).from_address(pb)
    for i in range(100):
        ia[i] = i * 3
    L.copyMemory(pb, pa, 400)
    assert list(ib) == [i * 3 for i in range(100)]
    L.deleteArray(a)
    L.deleteArray(b)
