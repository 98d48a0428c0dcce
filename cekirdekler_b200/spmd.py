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


def attach(cruncher, rank: int, world: int, device=None):
    """Put `cruncher` (one device) into SPMD mode as global device `rank` of `world`."""
    if world > 1:
        cruncher.setSpmd(rank, world, make_exchange(rank, world, device))
    return cruncher
