"""One process per GPU on real devices (world size 2, NCCL): the range split, the time exchange and — the one collective of
the path — the whole-array `read` input that crosses one rank's host link and reaches the other rank by
ncclBroadcast over NVLink.  The proof that the broadcast carried the data: on the call whose root is rank 0, rank 1's own
host copy of the positions is poisoned, and rank 1's slice of the result must still match the oracle."""
import os
import socket
import subprocess
import sys
from pathlib import Path

import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent

WORKER = r'''
import os, sys
import numpy as np
sys.path.insert(0, os.environ["CK_ROOT"])
import torch, torch.distributed as dist
import cekirdekler_b200 as ck
from cekirdekler_b200 import kernels, spmd
from cekirdekler_b200._native import shim
from oracle import pyoracle as po

rank, world, local = spmd.init_process_group_from_env("nccl")
dev = torch.device("cuda", local)
n, steps, wpb, L = 8192, 4, 4, 128
pos0 = np.zeros((n, 4), np.float32)
for c in range(3):
    pos0[:, c] = po.uniform(0xC0FFEE + 30 + c, n, -15, 15)
pos0[:, 3] = 1.0
params_h = np.array([1e-3, 1e-4, n, 0], np.float32)
p, v, o = pos0.copy().reshape(-1), np.zeros(4 * n, np.float32), np.zeros(4 * n, np.float32)
expected = []
for _ in range(steps):
    po.nbody3d_step(p, v, o, params_h)
    p, o = o, p
    expected.append(p.reshape(n, 4).copy())

cr = ck.ClNumberCruncher(ck.ClDevices([local]), kernels.source("nbody"))
assert cr.errorCode() == 0, cr.errorMessage()
spmd.attach(cr, rank, world, dev)
want_broadcast = os.environ.get("CK_SPMD_BROADCAST", "1") != "0"
assert (shim().ckSpmdWorld() == world) == want_broadcast, shim().ckGetLastError()
kfloats = ((n // 128 + 8) * ((n // 256 + 7) // 8) * 384 + 4096) // 1024
pa, pb, vel, params, scratch = (ck.ClFloatArray(4 * n), ck.ClFloatArray(4 * n), ck.ClFloatArray(4 * n), ck.ClFloatArray(8),
                                ck.ClFloatArray(kfloats * 1024))
pa.view[:] = pos0.reshape(-1)
params.view[:5] = [1e-3, 1e-4, n, wpb, kfloats]
params.readOnly = True
scratch.read, scratch.write = False, False
vel.partialRead, vel.read = True, False
for a in (pa, pb, vel):
    a.numberOfElementsPerWorkItem = 4 // wpb
src, dst = pa, pb
ok = True
for s in range(steps):
    src.read, src.partialRead, src.write = True, False, False
    dst.read, dst.partialRead, dst.write = False, False, True
    # whole-read arrays of this call, in argument order: src (root = 2s mod 2 = 0) and params (root 1).  With the
    # broadcast on, rank 1 never needs its own copy of src: poison it.
    if want_broadcast and rank == 1:
        src.view[:] = np.nan
    assert src.nextParam(vel, dst, params, scratch).compute(cr, 1, "nBody3Dw", n * wpb, L) == 0
    r, f = cr.globalRanges(1), cr.globalReferences(1)
    lo, cnt = f[rank] // wpb, r[rank] // wpb
    got = dst.view.reshape(n, 4)[lo:lo + cnt]
    exp = expected[s][lo:lo + cnt]
    rel = np.abs(got[:, :3] - exp[:, :3]) / np.maximum(np.abs(exp[:, :3]), 1e-3)
    ok = ok and cnt > 0 and bool(np.isfinite(got).all()) and float(rel.max()) <= 1e-4
    # the next step reads every body, and the split may move: assemble the full arrays on every rank (the caller's job
    # when each rank keeps its own host arrays)
    for arr in (dst, vel):
        full = torch.zeros(4 * n, dtype=torch.float32, device=dev)
        full[4 * lo:4 * (lo + cnt)] = torch.from_numpy(arr.view[4 * lo:4 * (lo + cnt)].copy()).to(dev)
        dist.all_reduce(full)
        arr.view[:] = full.cpu().numpy()
    src, dst = dst, src
flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
print("RESULT", rank, int(flag.item()), cr.globalRanges(1))
for a in (pa, pb, vel, params, scratch):
    a.dispose()
cr.dispose()
shim().ckSpmdFinalize()
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if int(flag.item()) == 1 else 1)
'''


WORKER_PEER = r"""
import os, sys
import numpy as np
sys.path.insert(0, os.environ["CK_ROOT"])
import torch, torch.distributed as dist
import cekirdekler_b200 as ck
from cekirdekler_b200 import kernels, spmd
from oracle import pyoracle as po

rank, world, local = spmd.init_process_group_from_env("nccl")
dev = torch.device("cuda", local)
n, steps, wpb, L = 8192, 6, 4, 128
pos0 = np.zeros((n, 4), np.float32)
for c in range(3):
    pos0[:, c] = po.uniform(0xC0FFEE + 40 + c, n, -15, 15)
pos0[:, 3] = 1.0
params_h = np.array([1e-3, 1e-4, n, 0], np.float32)
p, v, o = pos0.copy().reshape(-1), np.zeros(4 * n, np.float32), np.zeros(4 * n, np.float32)
for _ in range(steps):
    po.nbody3d_step(p, v, o, params_h)
    p, o = o, p
cr = ck.ClNumberCruncher(ck.ClDevices([local]), kernels.source("nbody"))
assert cr.errorCode() == 0, cr.errorMessage()
spmd.attach(cr, rank, world, dev)
kfloats = ((n // 128 + 8) * ((n // 256 + 7) // 8) * 384 + 4096) // 1024
pa, pb, vel, params, scratch = (ck.ClFloatArray(4 * n), ck.ClFloatArray(4 * n), ck.ClFloatArray(4 * n), ck.ClFloatArray(8),
                                ck.ClFloatArray(kfloats * 1024))
pa.view[:] = pos0.reshape(-1)
params.view[:5] = [1e-3, 1e-4, n, wpb, kfloats]
params.readOnly = True
scratch.read, scratch.write = False, False
for a in (pa, pb, vel):
    a.numberOfElementsPerWorkItem = 4 // wpb
tables = {id(pa): spmd.PeerTable(cr, rank, world, dev, pa, vel), id(pb): spmd.PeerTable(cr, rank, world, dev, pb, vel)}
assert all(t.ok for t in tables.values()), [t.error for t in tables.values()]
src, dst = pa, pb
for s in range(steps):
    last = s == steps - 1
    src.read, src.partialRead, src.write = (s == 0), False, False     # every rank uploads the positions once
    dst.read, dst.partialRead, dst.write = False, False, last
    vel.read, vel.partialRead, vel.write = False, False, last
    assert src.nextParam(vel, dst, params, scratch, tables[id(dst)].array).compute(cr, 1, "nBody3Dwp", n * wpb, L) == 0
    src, dst = dst, src
r, f = cr.globalRanges(1), cr.globalReferences(1)
lo, cnt = f[rank] // wpb, r[rank] // wpb
got, exp = src.view.reshape(n, 4)[lo:lo + cnt], p.reshape(n, 4)[lo:lo + cnt]
rel = np.abs(got[:, :3] - exp[:, :3]) / np.maximum(np.abs(exp[:, :3]), 1e-3)
vexp = v.reshape(n, 4)[lo:lo + cnt, :3]
ok = cnt > 0 and float(rel.max()) <= 1e-4 and float(np.abs(vel.view.reshape(n, 4)[lo:lo + cnt, :3] - vexp).max() / np.abs(vexp).max()) <= 1e-4
flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
print("RESULT", rank, int(flag.item()), r, float(rel.max()))
for t in tables.values():
    t.close()
for a in (pa, pb, vel, params, scratch):
    a.dispose()
cr.dispose()
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if int(flag.item()) == 1 else 1)
"""


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def run_two_ranks(worker, **extra_env):
    port = free_port()
    procs = []
    for rank in range(2):
        env = dict(os.environ, **extra_env)
        env.update( RANK=str(rank), WORLD_SIZE="2", LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
                   CK_ROOT=str(ROOT))
        procs.append(subprocess.Popen([sys.executable, "-c", worker], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    outs = []
    for p in procs:
        try:
            out, err = p.communicate(timeout=300)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        outs.append((p.returncode, out, err))
    for rc, out, err in outs:
        assert rc == 0, (out[-2000:], err[-4000:])
        assert "RESULT" in out


@pytest.mark.parametrize("broadcast", ["1", "0"], ids=["nccl-broadcast", "per-rank-upload"])
def test_two_ranks_on_two_gpus(gpus, broadcast):
    if gpus < 2:
        pytest.skip("needs two GPUs")
    # defaults are world >= 4 and arrays >= 16 MiB: force the broadcast on for this small case
    run_two_ranks(WORKER, CK_SPMD_BROADCAST=broadcast, CK_SPMD_BROADCAST_MIN_WORLD="2", CK_SPMD_BROADCAST_MIN_BYTES="0")


def test_two_ranks_resident_steps_with_peer_stores(gpus):
    """nBody3Dwp across PROCESSES: each rank's kernel stores its new positions and velocities into the other rank's buffers,
    mapped through CUDA IPC (spmd.PeerTable); positions are uploaded once and only the last step downloads.  The exchange of
    device times that opens every compute() is the barrier between steps."""
    if gpus < 2:
        pytest.skip("needs two GPUs")
    run_two_ranks(WORKER_PEER)
