// ext_ipc.cu — device buffers of OTHER PROCESSES, for kernels that store into every GPU's copy of an array over NVLink
// (nBody3Dwp) when each GPU has its own process (torchrun).  No counterpart in the reference, which is single-process.
//
//   ckIpcExport(device, ptr, out64)   64-byte handle of the allocation that `ptr` (a cudaMalloc'd buffer of this process,
//                                     e.g. ckBufferDevicePointer) starts; the launcher's own channel carries it to the
//                                     other ranks (cekirdekler_b200/spmd.py uses torch.distributed)
//   ckIpcOpen(device, handle64)       maps another process's allocation into this one; the returned device pointer is
//                                     valid for kernels and copies on `device` (peer access is enabled on demand)
//   ckIpcClose(device, ptr)           unmaps it; the owner must keep the buffer alive until every importer has closed it
#include "handles.h"
#include "../../../include/KutuphaneCL.h"

#include <cstring>
#include <string>

extern "C" {

int32_t ckIpcExport(int32_t device, void* devicePointer, void* out64) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    if (!devicePointer || !out64) {
        ck::set_error("ckIpcExport: null pointer");
        return 1;
    }
    if (!ck::cuda_ok(cudaSetDevice(device), "ckIpcExport: cudaSetDevice")) return 2;
    cudaIpcMemHandle_t h;
    if (!ck::cuda_ok(cudaIpcGetMemHandle(&h, devicePointer), "ckIpcExport: cudaIpcGetMemHandle")) return 3;
    memcpy(out64, &h, sizeof h);
    return 0;
}

void* ckIpcOpen(int32_t device, const void* handle64) {
    if (!handle64) {
        ck::set_error("ckIpcOpen: null handle");
        return nullptr;
    }
    if (!ck::cuda_ok(cudaSetDevice(device), "ckIpcOpen: cudaSetDevice")) return nullptr;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, sizeof h);
    void* p = nullptr;
    if (!ck::cuda_ok(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess), "ckIpcOpen: cudaIpcOpenMemHandle")) return nullptr;
    return p;
}

int32_t ckIpcClose(int32_t device, void* importedPointer) {
    if (!importedPointer) return 0;
    if (!ck::cuda_ok(cudaSetDevice(device), "ckIpcClose: cudaSetDevice")) return 2;
    return ck::cuda_ok(cudaIpcCloseMemHandle(importedPointer), "ckIpcClose: cudaIpcCloseMemHandle") ? 0 : 3;
}

}  // extern "C"
