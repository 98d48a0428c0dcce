"""ctypes loaders for the two native libraries.  There is no Python/CPU fallback: if a library is missing the
import of anything that computes fails loudly (build with ``python -m cekirdekler_b200._build``)."""
from __future__ import annotations

import ctypes as C
from pathlib import Path

_LIB = Path(__file__).resolve().parent / "lib"
_shim = None
_host = None


class NativeLibraryMissing(ImportError):
    pass


def _load(name: str) -> C.CDLL:
    path = _LIB / name
    if not path.exists():
        raise NativeLibraryMissing(
            f"{path} is not built. Run `python -c 'import __graft_entry__ as g; g.build()'` (needs nvcc); "
            "this package has no CPU fallback."
        )
    return C.CDLL(str(path), mode=C.RTLD_GLOBAL)


def shim() -> C.CDLL:
    """libKutuphaneCL.so — the 81-symbol C ABI (include/KutuphaneCL.h)."""
    global _shim
    if _shim is None:
        L = _load("libKutuphaneCL.so")
        vp, i32, u64, i64, dbl, cp = C.c_void_p, C.c_int32, C.c_uint64, C.c_int64, C.c_double, C.c_char_p
        sig = {
            "platformList": (vp, []), "numberOfPlatforms": (i32, [vp]), "deletePlatformList": (None, [vp]),
            "createPlatform": (vp, [vp, i32]), "deletePlatform": (None, [vp]), "getPlatformInfo": (vp, [vp]),
            "getPlatformNameString": (vp, [vp]), "getPlatformVendorNameString": (vp, [vp]),
            "CODE_CPU": (i32, []), "CODE_GPU": (i32, []), "CODE_ACC": (i32, []),
            "numberOfCpusInPlatform": (i32, [vp]), "numberOfGpusInPlatform": (i32, [vp]),
            "numberOfAcceleratorsInPlatform": (i32, [vp]),
            "createDevice": (vp, [vp, i32, i32]), "createDeviceAsPartition": (vp, [vp, i32, i32, i32]),
            "deleteDevice": (None, [vp]), "getDeviceName": (None, [vp, vp]), "getDeviceVendorName": (None, [vp, vp]),
            "deviceGDDR": (i32, [vp]), "deviceComputeUnits": (i32, [vp]), "deviceMemSize": (u64, [vp]),
            "createContext": (vp, [vp, vp]), "deleteContext": (None, [vp]),
            "createCommandQueue": (vp, [vp, vp, i32]), "createCommandQueue2": (vp, [vp, vp, i32, i32]),
            "deleteCommandQueue": (None, [vp]), "addMarkerToCommandQueue": (None, [vp]),
            "getMarkerCounterOfCommandQueue": (i32, [vp]), "resetMarkerCounterOfCommandQueue": (None, [vp]),
            "createEvent": (vp, []), "deleteEvent": (None, [vp]), "createEventArr": (vp, [i32]),
            "addToEventArr": (None, [vp, vp, i32]), "deleteEventArr": (None, [vp]),
            "createUserEvent": (vp, [vp]), "deleteUserEvent": (None, [vp]), "triggerUserEvent": (None, [vp]),
            "addUserEvent": (None, [vp, vp]), "decrementUserEvent": (None, [vp, vp]), "incrementUserEvent": (None, [vp]),
            "createString": (vp, []), "writeToString": (vp, [vp, cp]), "deleteString": (None, [vp]),
            "readFromString": (cp, [vp]), "jsonStringCallBack": (None, [vp]),
            "createProgram": (vp, [vp, vp, vp]), "deleteProgram": (None, [vp]), "getProgramErr": (i32, [vp]),
            "readProgramErrString": (vp, [vp]),
            "createKernel": (vp, [vp, vp]), "deleteKernel": (None, [vp]), "getKernelErr": (i32, [vp]),
            "setKernelArgument": (None, [vp, vp, i32]), "setKernelArguments": (None, [vp, vp, i32]),
            "createNdRange": (vp, [i32]), "deleteNdRange": (None, [vp]),
            "compute": (i32, [vp, vp, vp, vp, vp]), "computeRepeated": (i32, [vp, vp, vp, vp, vp, i32]),
            "computeRepeatedWithSyncKernel": (i32, [vp, vp, vp, vp, vp, i32, vp, vp]),
            "computeEvent": (i32, [vp, vp, vp, vp, vp, vp, vp]),
            "finish": (None, [vp]), "flush": (None, [vp]), "wait2": (None, [vp, vp]), "wait3": (None, [vp, vp, vp]),
            "waitN": (None, [C.POINTER(vp), vp, i32]),
            "createBuffer": (vp, [vp, i32, i32, i32, vp, i32, i32, i32]), "deleteBuffer": (None, [vp]),
            "writeToBuffer": (vp, [vp, vp, vp]), "writeToBufferEvent": (vp, [vp, vp, vp, vp, vp]),
            "writeToBufferRanged": (vp, [vp, vp, i32, i32, vp]),
            "writeToBufferRangedEvent": (vp, [vp, vp, i32, i32, vp, vp, vp]),
            "readFromBuffer": (vp, [vp, vp, vp]), "readFromBufferEvent": (vp, [vp, vp, vp, vp, vp]),
            "readFromBufferRanged": (vp, [vp, vp, i32, i32, vp]),
            "readFromBufferRangedEvent": (vp, [vp, vp, i32, i32, vp, vp, vp]),
            "sizeOf": (None, [C.POINTER(i32)]), "createArray": (vp, [i32, i32, i32]), "alignedArrHead": (vp, [vp]),
            "deleteArray": (None, [vp]), "copyMemory": (None, [vp, vp, C.c_uint32]),
            # extensions
            "ckGetLastError": (cp, []), "ckLaunchCount": (u64, []), "ckDeviceCount": (i32, []),
            "ckCreateTimer": (vp, [vp]), "ckDeleteTimer": (None, [vp]), "ckTimerStart": (None, [vp, vp]),
            "ckTimerStop": (None, [vp, vp]), "ckTimerElapsedMs": (dbl, [vp]),
            "ckBufferDevicePointer": (vp, [vp]), "ckBufferBytes": (i64, [vp]), "ckBufferIsZeroCopy": (i32, [vp]),
            "ckQueueStream": (vp, [vp]), "ckContextDeviceOrdinal": (i32, [vp]),
            "ckCopyBufferPeer": (i32, [vp, vp, vp, i32, i32, vp, vp]), "ckPeerAccess": (i32, [i32, i32]),
            "ckQueueWaitEvent": (None, [vp, vp]), "ckQueueRecordEvent": (None, [vp, vp]),
            "ckEventSynchronize": (None, [vp]),
            "ckHostRegister": (i32, [vp, u64]), "ckHostUnregister": (i32, [vp]), "ckZeroBuffer": (i32, [vp, vp]),
            "ckBroadcastUpload": (i32, [C.POINTER(vp), C.POINTER(vp), i32, vp]),
            "ckBroadcastCounts": (None, [C.POINTER(u64), C.POINTER(u64)]),
            "ckNBody3DStepNoAlloc": (i32, [i32, vp, vp, vp, vp, i32, i32, i32, C.c_float, C.c_float]),
            "ckIpcExport": (i32, [i32, vp, vp]), "ckIpcOpen": (vp, [i32, vp]), "ckIpcClose": (i32, [i32, vp]),
            "ckSpmdUniqueId": (i32, [vp]), "ckSpmdInit": (i32, [vp, i32, i32, i32]), "ckSpmdWorld": (i32, []),
            "ckSpmdBroadcast": (i32, [vp, vp, i32]), "ckSpmdFinalize": (None, []),
        }
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        L._ck_signatures = sig
        _shim = L
    return _shim


ABI_SYMBOLS = None  # filled lazily by abi_symbols()


def abi_symbols() -> list:
    """Names declared in include/KutuphaneCL.h (parsed from the header, so the test follows the header)."""
    import re

    hdr = (Path(__file__).resolve().parent.parent / "include" / "KutuphaneCL.h").read_text()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = re.findall(r"\b([A-Za-z_][A-Za-z0-9_]*)\s*\(", hdr)
    skip = {"defined", "int", "void"}
    out = []
    for n in names:
        if n not in skip and n not in out:
            out.append(n)
    return out


def host() -> C.CDLL:
    """libcekirdekler_host.so — C++ host layer (include/cekirdekler_host.h)."""
    global _host
    if _host is None:
        shim()  # make sure libKutuphaneCL.so is resident first (the host library links against it)
        L = _load("libcekirdekler_host.so")
        vp, i32, u64, i64, dbl, cp = C.c_void_p, C.c_int32, C.c_uint64, C.c_int64, C.c_double, C.c_char_p
        pi32, pdbl = C.POINTER(i32), C.POINTER(dbl)
        L.EXCHANGE_FN = C.CFUNCTYPE(None, pdbl, i32, vp)
        sig = {
            "ckh_gpu_count": (i32, []), "ckh_gpu_name": (cp, [i32]), "ckh_gpu_compute_units": (i32, [i32]),
            "ckh_gpu_memory_size": (u64, [i32]),
            "ckh_array_create": (vp, [i32, i64, i32]), "ckh_array_wrap": (vp, [vp, i32, i64]),
            "ckh_array_destroy": (None, [vp]), "ckh_array_data": (vp, [vp]), "ckh_array_length": (i64, [vp]),
            "ckh_array_is_fast": (i32, [vp]), "ckh_array_set": (None, [vp, i32, i32]), "ckh_array_get": (i32, [vp, i32]),
            "ckh_array_flag_string": (cp, [vp]),
            "ckh_cruncher_create": (vp, [pi32, i32, cp, i32, i32, i32]), "ckh_cruncher_destroy": (None, [vp]),
            "ckh_cruncher_error_code": (i32, [vp]), "ckh_cruncher_error_message": (cp, [vp]),
            "ckh_cruncher_errors_happened": (i32, [vp]), "ckh_cruncher_num_devices": (i32, [vp]),
            "ckh_cruncher_device_name": (cp, [vp, i32]), "ckh_cruncher_num_kernels": (i32, [cp]),
            "ckh_cruncher_kernel_name": (cp, [cp, i32]),
            "ckh_cruncher_set": (None, [vp, i32, i32]), "ckh_cruncher_get": (i32, [vp, i32]),
            "ckh_cruncher_set_repeat_kernel_name": (None, [vp, cp]),
            "ckh_cruncher_markers_remaining": (i32, [vp]), "ckh_cruncher_markers_reached": (i32, [vp]),
            "ckh_cruncher_flush_last_used_queue": (None, [vp]), "ckh_cruncher_sync": (None, [vp]),
            "ckh_cruncher_performance_report": (cp, [vp, i32]),
            "ckh_cruncher_performance_history": (i32, [vp, i32, pdbl, i32]),
            "ckh_cruncher_buffer_pointer": (u64, [vp, vp, i32]),
            "ckh_cruncher_ranges": (i32, [vp, i32, pi32, i32]), "ckh_cruncher_references": (i32, [vp, i32, pi32, i32]),
            "ckh_cruncher_benchmarks": (i32, [vp, i32, pdbl, i32]),
            "ckh_cruncher_normalized_compute_powers": (i32, [vp, pdbl, i32]),
            "ckh_cruncher_normalized_global_ranges": (i32, [vp, i32, pdbl, i32]),
            "ckh_cruncher_reset_balancer": (None, [vp, i32]),
            "ckh_cruncher_set_spmd": (None, [vp, i32, i32, L.EXCHANGE_FN, vp]),
            "ckh_compute": (i32, [vp, C.POINTER(vp), i32, i32, cp, i32, i32, i32, i32, i32, i32, C.POINTER(cp), pi32]),
            "ckh_balancer_create": (vp, [i32, i32]), "ckh_balancer_destroy": (None, [vp]),
            "ckh_balancer_step": (None, [vp, pdbl, i32, i32, pi32, i32, i32, i32]),
            "ckh_load_balance": (None, [pdbl, i32, pdbl, i32, pi32, i32, i32]),
            "ckh_stage_create": (vp, [i32]), "ckh_stage_destroy": (None, [vp]),
            "ckh_stage_add_devices": (None, [vp, pi32, i32]),
            "ckh_stage_add_kernels": (None, [vp, cp, cp, pi32, pi32, i32]),
            "ckh_stage_initializer_kernel": (None, [vp, cp, pi32, pi32, i32]),
            "ckh_stage_add_buffer": (None, [vp, i32, i32, i64, i32]),
            "ckh_stage_prepend_to": (None, [vp, vp]), "ckh_stage_append_to": (None, [vp, vp]),
            "ckh_stage_set_enqueue_mode": (None, [vp, i32]), "ckh_stage_make_pipeline": (vp, [vp, i32]),
            "ckh_pipeline_destroy": (None, [vp]),
            "ckh_pipeline_push": (i32, [vp, C.POINTER(vp), i32, C.POINTER(vp), i32]),
            "ckh_pipeline_error_code": (i32, [vp]), "ckh_pipeline_error_message": (cp, [vp]),
            "ckh_pipeline_forwarding": (i32, [vp]), "ckh_pipeline_counter": (i32, [vp]),
            "ckh_dparray_wrap": (vp, [i32, vp]), "ckh_dparray_create": (vp, [i32, i32, i64, i32]),
            "ckh_dparray_destroy": (None, [vp]),
            "ckh_dpstage_create": (vp, [cp, i32, i32]), "ckh_dpstage_destroy": (None, [vp]),
            "ckh_dpstage_bind_array": (None, [vp, vp]), "ckh_dpstage_stop_transmission": (None, [vp, i32]),
            "ckh_dpipeline_create": (vp, [i32, cp, i32]), "ckh_dpipeline_destroy": (None, [vp]),
            "ckh_dpipeline_add_stage": (None, [vp, vp]), "ckh_dpipeline_feed": (None, [vp]),
            "ckh_dpipeline_feed_async_begin": (None, [vp]), "ckh_dpipeline_feed_async_end": (None, [vp]),
            "ckh_dpipeline_set_serial": (None, [vp, i32]), "ckh_dpipeline_error_code": (i32, [vp]),
            "ckh_dpipeline_error_message": (cp, [vp]), "ckh_dpipeline_latency": (i32, [vp]),
            "ckh_task_create": (vp, [C.POINTER(vp), i32, C.POINTER(cp), pi32, i32, cp, i32, i32, i32, i32, i32, i32]),
            "ckh_task_destroy": (None, [vp]), "ckh_task_set_type": (None, [vp, i32]), "ckh_task_get_type": (i32, [vp]),
            "ckh_task_set_callback": (None, [vp, vp, vp]), "ckh_task_compute": (i32, [vp, vp]),
            "ckh_taskpool_create": (vp, []), "ckh_taskpool_destroy": (None, [vp]), "ckh_taskpool_feed": (None, [vp, vp]),
            "ckh_taskpool_reset": (None, [vp]), "ckh_taskpool_remaining": (i32, [vp]),
            "ckh_devicepool_create": (vp, [i32, cp, i32, i32]), "ckh_devicepool_destroy": (None, [vp]),
            "ckh_devicepool_add_device": (None, [vp, i32]), "ckh_devicepool_enqueue": (None, [vp, vp]),
            "ckh_devicepool_finish": (dbl, [vp]), "ckh_devicepool_dispose": (None, [vp]),
            "ckh_devicepool_error_code": (i32, [vp]), "ckh_devicepool_error_message": (cp, [vp]),
            "ckh_devicepool_tasks_per_device": (i32, [vp, pi32, i32]),
        }
        L.CALLBACK_FN = C.CFUNCTYPE(None, vp)
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        L._ckh_signatures = sig
        _host = L
    return _host
