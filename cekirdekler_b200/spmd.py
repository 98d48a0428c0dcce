"""One process per GPU (torchrun): the same range split as the in-process multi-device path, without a data-path
collective.

Every rank owns one GPU and one single-device ClNumberCruncher.  The only thing ranks exchange is the vector of
per-device times of the previous call (world doubles: through a shared-memory file when all ranks share a host, else
all-gathered through torch.distributed); each rank then runs the same deterministic balancer (Cores.cs:544-613 restated in
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


_exchange_count = 0


class _SharedSlots:
    """`world` (sequence, value) slots in a POSIX shared-memory file, two banks deep: the ranks of one host exchange one
    double per compute() call through plain memory instead of a collective.

    Why not the process group: every compute() of a one-process-per-GPU job opens with this exchange (the balancer needs
    all devices' times of the previous call).  Through NCCL that is a kernel launch, a device synchronisation and a
    device-to-host copy per call, about 0.15 ms on eight B200s, as much as a sixth of an n-body step.  Here a rank stores
    its value, then its call number (x86 keeps the two stores in order), and reads the others' once their call numbers match.
    Call k uses bank k % 2; a rank reaches call k + 2 only after every rank has posted call k + 1, i.e. after every rank
    has finished reading call k, so two banks are enough."""

    def __init__(self, rank: int, world: int, path: str, create: bool):
        import mmap

        import numpy as np

        size = 2 * world * 64
        flags = os.O_RDWR | (os.O_CREAT | os.O_EXCL if create else 0)
        fd = os.open(path, flags, 0o600)
        try:
            if create:
                os.ftruncate(fd, size)
            self._mm = mmap.mmap(fd, size)
        finally:
            os.close(fd)
        self._seq = np.frombuffer(self._mm, dtype=np.int64).reshape(2, world, 8)[:, :, 0]
        self._val = np.frombuffer(self._mm, dtype=np.float64).reshape(2, world, 8)[:, :, 1]
        self.rank, self.world, self.calls = rank, world, 0

    def exchange(self, value: float, timeout_s: float = 300.0) -> List[float]:
        import time

        self.calls += 1
        k, bank = self.calls, self.calls % 2
        self._val[bank, self.rank] = value
        self._seq[bank, self.rank] = k
        seq, spins, t0 = self._seq[bank], 0, None
        while not (seq == k).all():
            spins += 1
            if spins > 20000:                       # the others are far behind (first call, a slower share): stop burning the core
                if t0 is None:
                    t0 = time.monotonic()
                elif time.monotonic() - t0 > timeout_s:
                    raise RuntimeError(f"spmd exchange: rank {self.rank} waited {timeout_s:.0f} s for call {k} of the other ranks")
                time.sleep(20e-6)
        return self._val[bank].tolist()


def make_exchange(rank: int, world: int, device=None) -> Callable[[List[float]], List[float]]:
    """all-gather of one double per rank: in → list of `world` times with entry `rank` valid; out → all entries valid.

    When every rank of the job runs on this host (torchrun's LOCAL_WORLD_SIZE == WORLD_SIZE, the only layout the 8-GPU box
    has) the values travel through a shared-memory file (_SharedSlots) that the ranks agree on with ONE collective here, at
    set-up; CK_SPMD_EXCHANGE=collective, or ranks on several hosts, keep the per-call all-gather of the process group."""
    import torch
    import torch.distributed as dist

    global _exchange_count
    slots = None
    one_host = int(os.environ.get("LOCAL_WORLD_SIZE", str(world))) == world
    if world > 1 and dist.is_initialized() and one_host and os.environ.get("CK_SPMD_EXCHANGE", "shm") != "collective":
        _exchange_count += 1
        ident = torch.tensor([os.getpid() if rank == 0 else 0], dtype=torch.int64, device=device)
        dist.broadcast(ident, src=0)                                   # every rank makes its exchanges in the same order
        path = f"/dev/shm/ck_spmd_{int(ident.item())}_{_exchange_count}"
        ok = 1
        try:
            if rank == 0:
                slots = _SharedSlots(rank, world, path, create=True)
            dist.barrier()
            if rank != 0:
                slots = _SharedSlots(rank, world, path, create=False)
        except OSError:
            ok = 0
        flag = torch.tensor([ok], dtype=torch.int32, device=device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)                    # also: everybody has mapped the file by now
        if rank == 0:
            try:
                os.unlink(path)                                        # the mappings keep it alive; nothing is left behind
            except OSError:
                pass
        if not int(flag.item()):
            slots = None

    def exchange(times: List[float]) -> List[float]:
        if world <= 1 or not dist.is_initialized():
            return times
        if slots is not None:
            return slots.exchange(times[rank])
        mine = torch.tensor([times[rank]], dtype=torch.float64, device=device)
        out = torch.empty(world, dtype=torch.float64, device=device)
        dist.all_gather_into_tensor(out, mine)
        return out.cpu().tolist()

    exchange.uses_shared_memory = slots is not None
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
