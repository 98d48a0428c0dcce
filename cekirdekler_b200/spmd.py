"""One process per GPU (torchrun): the same range split as the in-process multi-device path, without a data-path
collective.

Every rank owns one GPU and one single-device ClNumberCruncher.  The only thing ranks exchange is the vector of
per-device times of the previous call (world doubles, all-gathered through torch.distributed — NCCL on GPU boxes,
gloo in the CPU tests); each rank then runs the same deterministic balancer (Cores.cs:544-613 restated in
csrc/host/balance.cpp) and so arrives at the same split, of which it executes its own slice [ref_r, ref_r + range_r).
torch is plumbing here (process group, device selection), not part of the compute path.
"""
from __future__ import annotations

import os
from typing import Callable, List


def init_process_group_from_env(backend: str | None = None):
    """Join the torchrun rendezvous described by RANK/WORLD_SIZE/MASTER_ADDR/MASTER_PORT; returns (rank, world, local_rank)."""
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, rank=rank, world_size=world, device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend, rank=rank, world_size=world)
    return rank, world, local


def make_exchange(rank: int, world: int, device=None) -> Callable[[List[float]], List[float]]:
    """all-gather of one double per rank: in → list of `world` times with entry `rank` valid; out → all entries valid."""
    import torch
    import torch.distributed as dist

    def exchange(times: List[float]) -> List[float]:
        if world <= 1 or not dist.is_initialized():
            return times
        mine = torch.tensor([times[rank]], dtype=torch.float64, device=device)
        out = torch.empty(world, dtype=torch.float64, device=device)
        dist.all_gather_into_tensor(out, mine)
        return out.cpu().tolist()

    return exchange


_broadcast_state = {"tried": False, "ok": False}


def init_rank_broadcast(rank: int, world: int, device_ordinal: int, device=None) -> bool:
    """Collective, once per process: builds the shim's NCCL communicator over the ranks (ckSpmdInit), which lets
    Cores.compute send a whole-array `read` input over ONE rank's host link and fan it out over NVLink
    instead of uploading it on every rank.  Rank 0 draws the NCCL id; torch.distributed carries its 128 bytes to the other
    ranks.  Returns whether every rank has the communicator; without NCCL or GPUs (gloo tests) it returns False and the
    ranks keep uploading for themselves.  CK_SPMD_BROADCAST=0 switches it off."""
    import ctypes

    import torch
    import torch.distributed as dist

    from ._native import shim

    st = _broadcast_state
    if st["tried"] or world <= 1 or not dist.is_initialized():
        return st["ok"]
    st["tried"] = True
    if os.environ.get("CK_SPMD_BROADCAST", "1") == "0" or os.environ.get("CK_BROADCAST") == "p2p":
        return False
    L = shim()
    buf = (ctypes.c_ubyte * 128)()
    have = 1
    if rank == 0:
        have = 1 if (L.ckDeviceCount() > 0 and L.ckSpmdUniqueId(buf) == 0) else 0
    t = torch.tensor([have] + list(buf), dtype=torch.int32, device=device)
    dist.broadcast(t, src=0)
    vals = t.cpu().tolist()
    ok = 0
    if vals[0] == 1 and L.ckDeviceCount() > 0:
        ident = (ctypes.c_ubyte * 128)(*vals[1:])
        ok = 1 if L.ckSpmdInit(ident, rank, world, device_ordinal) == 0 else 0
    elif vals[0] == 1:
        ok = 0
    # every rank must have joined, or nobody uses it
    if vals[0] == 1:
        flag = torch.tensor([ok], dtype=torch.int32, device=device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        ok = int(flag.item())
        if not ok:
            L.ckSpmdFinalize()
    st["ok"] = bool(ok)
    return st["ok"]


def attach(cruncher, rank: int, world: int, device=None):
    """Put `cruncher` (one device) into SPMD mode as global device `rank` of `world`."""
    if world > 1:
        cruncher.setSpmd(rank, world, make_exchange(rank, world, device))
        if device is not None and getattr(device, "type", "cpu") == "cuda":
            init_rank_broadcast(rank, world, cruncher.deviceOrdinals[0] if hasattr(cruncher, "deviceOrdinals") else device.index, device)
    return cruncher


class PeerTable:
    """The one-process-per-GPU form of ClNumberCruncher.peerTable: `array` is a whole-`read` ClArray holding
    { world, target0 on rank 0..world-1, target1 on rank 0..world-1, ... } as device addresses valid in THIS process — the
    own buffers directly, the other ranks' buffers through CUDA IPC (handles all-gathered over torch.distributed).
    Collective: every rank constructs it with the same targets in the same order, then checks `ok` (the same on every
    rank; `error` says what failed locally) before using `array`.  close() is collective too (a barrier first, so that
    nobody unmaps or frees a buffer another rank's kernel may still be writing)."""

    def __init__(self, cruncher, rank: int, world: int, device, *targets):
        import ctypes

        import numpy as np
        import torch
        import torch.distributed as dist

        from ._native import shim
        from .api import ClArray

        # Nothing below raises between the collectives: a rank that fails keeps taking part and reports through `ok`
        # (the minimum over ranks), so that either every rank uses the table or none does.
        self._L, self._dev, self._opened = shim(), device.index, []
        self.error = ""
        L, h = self._L, cruncher._h
        own = []
        blob = torch.zeros(len(targets) * 64, dtype=torch.uint8)
        for t, target in enumerate(targets):
            p = h.ckh_cruncher_buffer_pointer(cruncher._handle, target._handle, 0)
            own.append(p)
            raw = (ctypes.c_ubyte * 64)()
            if not p or L.ckIpcExport(self._dev, ctypes.c_void_p(p), raw) != 0:
                self.error = self.error or ("export of target %d: %s" % (t, L.ckGetLastError().decode()))
            blob[t * 64:(t + 1) * 64] = torch.tensor(list(raw), dtype=torch.uint8)
        gathered = torch.zeros(world * blob.numel(), dtype=torch.uint8, device=device)
        dist.all_gather_into_tensor(gathered, blob.to(device))
        handles = gathered.cpu().numpy().reshape(world, len(targets), 64)
        self.array = ClArray(1 + world * max(len(targets), 1), np.int64)
        self.array.view[0] = world
        for t in range(len(targets)):
            for r in range(world):
                if r == rank:
                    p = own[t]
                elif self.error:
                    p = 0
                else:
                    raw = (ctypes.c_ubyte * 64)(*handles[r, t].tolist())
                    p = L.ckIpcOpen(self._dev, raw) or 0
                    if p:
                        self._opened.append(p)
                    else:
                        self.error = "open of rank %d target %d: %s" % (r, t, L.ckGetLastError().decode())
                self.array.view[1 + t * world + r] = np.array([p], dtype=np.uint64).view(np.int64)[0]
        self.array.readOnly = True
        flag = torch.tensor([0 if self.error else 1], dtype=torch.int32, device=device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        self.ok = bool(flag.item())

    def close(self):
        import torch.distributed as dist

        if dist.is_initialized():
            dist.barrier()
        for p in self._opened:
            self._L.ckIpcClose(self._dev, p)
        self._opened = []
        if dist.is_initialized():
            dist.barrier()                     # owners free their buffers only after every importer has unmapped them
        self.array.dispose()


def prefer_host_memory_near_gpu(local_rank: int) -> dict:
    """Make this process allocate its host memory (the pinned arrays above all) on the NUMA node the GPU hangs off.

    With one process per GPU every rank streams its slice over its own host link; when the pinned pages sit on the other
    socket each transfer also crosses the inter-socket link, which the ranks then share.  Looks the GPU's PCI address up
    through NVML, reads its node from sysfs and sets the task's memory policy to MPOL_PREFERRED for that node
    (set_mempolicy(2); threads stay free to run anywhere, so the CPU baseline still gets every core).  Call it before the
    arrays are created.  Returns what it did; never raises (single-node hosts and containers without sysfs are fine)."""
    info = {"numa_node": None, "applied": False}
    try:
        import ctypes
        import pynvml

        pynvml.nvmlInit()
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(local_rank)).busId
        bus = (bus.decode() if isinstance(bus, bytes) else bus).lower()
        if len(bus.split(":")[0]) == 8:
            bus = bus[4:]                      # NVML prints an 8-digit domain, sysfs a 4-digit one
        with open(f"/sys/bus/pci/devices/{bus}/numa_node") as f:
            node = int(f.read().strip())
        info["numa_node"] = node
        if node < 0 or not os.path.isdir(f"/sys/devices/system/node/node{node}"):
            return info
        MPOL_PREFERRED, SYS_set_mempolicy = 1, 238          # x86-64
        mask = (ctypes.c_ulong * 16)()
        mask[node // 64] = 1 << (node % 64)
        libc = ctypes.CDLL(None, use_errno=True)
        rc = libc.syscall(SYS_set_mempolicy, MPOL_PREFERRED, mask, 16 * 64)
        info["applied"] = rc == 0
        if rc != 0:
            info["errno"] = ctypes.get_errno()
    except Exception as e:  # no NVML, no sysfs, seccomp: leave the default policy
        info["error"] = f"{type(e).__name__}: {e}"
    return info
