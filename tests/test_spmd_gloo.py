"""world_size-2 gloo run on CPU of the one-process-per-GPU host logic: ranks all-gather their device times, run the same
balancer, and must agree on a split whose slices partition the range (no data-path collective)."""
import os
import socket
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent

WORKER = r'''
import ctypes as C, json, os, sys
sys.path.insert(0, os.environ["CK_ROOT"])
from cekirdekler_b200 import spmd
from cekirdekler_b200._native import host
rank, world, _ = spmd.init_process_group_from_env("gloo")
ex = spmd.make_exchange(rank, world)
h = host()
b = h.ckh_balancer_create(world, int(os.environ.get('CK_TEST_BALANCER_MODE', '1')))
ranges = (C.c_int32 * world)()
total, local = 1 << 20, 256
speed = [1.0, 2.5][:world]                  # rank 1's "device" is 2.5x faster
trace = []
mine = 0.0
for call in range(12):
    times = [0.0] * world
    times[rank] = mine
    times = ex(times)                        # the only exchange between ranks: world doubles
    t = (C.c_double * world)(*times)
    h.ckh_balancer_step(b, t, 1, total, ranges, 0, 4, local)
    r = list(ranges)
    refs = [sum(r[:i]) for i in range(world)]
    mine = r[rank] / speed[rank] * 1e-3      # synthetic device time of my slice
    trace.append({"ranges": r, "refs": refs, "times": times})
print("TRACE" + json.dumps(trace))
import torch.distributed as dist
dist.barrier()
dist.destroy_process_group()
'''


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


import pytest


@pytest.mark.parametrize("mode", [1, 2], ids=["adaptive", "profile"])
def test_two_ranks_agree_on_the_split(mode):
    import json

    port = free_port()
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE="2", LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), CK_ROOT=str(ROOT), OMP_NUM_THREADS="1", CK_TEST_BALANCER_MODE=str(mode))
        procs.append(subprocess.Popen([sys.executable, "-c", WORKER], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    outs = []
    for p in procs:
        out, err = p.communicate(timeout=240)
        assert p.returncode == 0, err[-2000:]
        outs.append(json.loads([l for l in out.splitlines() if l.startswith("TRACE")][0][5:]))
    a, b = outs
    assert a == b                                               # identical decisions on both ranks
    for step in a:
        r, refs = step["ranges"], step["refs"]
        assert sum(r) == 1 << 20 and all(x % 256 == 0 and x > 0 for x in r)
        assert refs == [0, r[0]]                                # slices tile the range with no gap or overlap
    last = a[-1]["ranges"]
    assert abs(last[1] / last[0] - 2.5) < 0.15                  # converged to the speed ratio
    # from the second call on, both ranks saw both times
    assert all(t > 0 for t in a[2]["times"])


def test_shared_memory_slots_exchange_many_calls_between_threads(tmp_path):
    """spmd._SharedSlots on its own (no process group): three "ranks" on three threads exchange 200 values; every rank
    must see, at every call, exactly the values the others posted for THAT call (two banks are enough: nobody can be more
    than one call ahead of the slowest reader)."""
    import threading

    from cekirdekler_b200.spmd import _SharedSlots

    world, calls = 3, 200      # threads share the GIL, so every hand-over costs a switch interval; ranks are processes
    path = str(tmp_path / "slots")
    slots = [_SharedSlots(0, world, path, create=True)] + [_SharedSlots(r, world, path, create=False) for r in range(1, world)]
    seen = [[] for _ in range(world)]

    def body(r):
        for k in range(calls):
            seen[r].append(slots[r].exchange(1000.0 * k + r, timeout_s=60.0))

    threads = [threading.Thread(target=body, args=(r,)) for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for r in range(world):
        assert len(seen[r]) == calls
        for k in (0, 1, 2, 77, calls - 1):
            assert seen[r][k] == [1000.0 * k + q for q in range(world)]
    assert all(seen[0][k] == seen[1][k] == seen[2][k] for k in range(calls))
