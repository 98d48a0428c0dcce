"""Task pool / device pool (ClPipeline.cs:3241-5078): frozen compute() calls scheduled greedily over the devices.
Every task's result is checked exactly; the control bits (global synchronisation, device select, serial mode, broadcast)
are checked through effects that only hold if the ordering they promise was kept."""
import threading

import numpy as np
import pytest

import cekirdekler_b200 as ck
from cekirdekler_b200.pool import ClDevicePool, ClDevicePoolType, ClTaskPool, ClTaskType, task

pytestmark = pytest.mark.gpu

SRC = """
__kernel void scale(__global int* a, __global int* k, __global int* out) { int i = get_global_id(0); out[i] = a[i] * k[0]; }
__kernel void addOne(__global int* a) { int i = get_global_id(0); a[i] = a[i] + 1; }
__kernel void setTable(__global int* table) { int i = get_global_id(0); table[i] = 7; }
__kernel void useTable(__global int* table, __global int* out) { int i = get_global_id(0); out[i] = table[i] + i; }
"""
N = 1 << 14


def devices():
    return ck.ClPlatforms.all().gpus()


@pytest.mark.parametrize("fine", [False, True], ids=["blocking", "fineGrained"])
@pytest.mark.parametrize("ptype", [ClDevicePoolType.DEVICE_COMPUTE_AT_WILL, ClDevicePoolType.DEVICE_ROUND_ROBIN])
def test_independent_tasks_all_complete_exactly(gpus, fine, ptype):
    pool = ClDevicePool(ptype, SRC, fine, 3)
    pool.addDevice(devices())
    assert pool.errorCode() == 0, pool.errorMessage()
    tp = ClTaskPool()
    arrays, done = [], []
    lock = threading.Lock()
    for t in range(48):
        a, k, out = ck.ClArray(N, np.int32), ck.ClArray(1, np.int32), ck.ClArray(N, np.int32)
        a.view[:] = np.arange(N) + t
        k[0] = t + 2
        a.readOnly, k.readOnly, out.writeOnly = True, True, True
        tk = task(a.nextParam(k, out), 100 + t, "scale", N, 256)

        def cb(t=t):
            with lock:
                done.append(t)
        tk.setCallBack(cb)
        tp.feed(tk)
        arrays.append((a, k, out))
    assert tp.remainingTaskGroupsOrTasks() == 48
    pool.enqueueTaskPool(tp)
    ms = pool.finish()
    assert ms > 0
    for t, (a, k, out) in enumerate(arrays):
        assert np.array_equal(out.view, (np.arange(N) + t) * (t + 2)), t
    assert sorted(done) == list(range(48))            # every callback ran exactly once, after its task completed
    per = pool.tasksPerDevice()
    assert sum(per) == 48 and len(per) == gpus
    if ptype == ClDevicePoolType.DEVICE_ROUND_ROBIN and gpus > 1:
        assert max(per) - min(per) <= 1
    pool.dispose()


def test_global_sync_orders_dependent_tasks(gpus):
    """Tasks on one array with GLOBAL_SYNCHRONIZATION_FIRST between them behave like a chain: a[i] ends at +5."""
    pool = ClDevicePool(ClDevicePoolType.DEVICE_COMPUTE_AT_WILL, SRC, True, 3)
    pool.addDevice(devices())
    a = ck.ClArray(N, np.int32)
    a.view[:] = np.arange(N)
    tp = ClTaskPool()
    for step in range(5):
        tk = task(a, 7, "addOne", N, 256)              # read whole + write slice: a true read-modify-write of the host array
        tk.type = ClTaskType.TASK_MESSAGE_GLOBAL_SYNCHRONIZATION_FIRST
        tp.feed(tk)
    pool.enqueueTaskPool(tp)
    pool.finish()
    assert np.array_equal(a.view, np.arange(N) + 5)
    pool.dispose()


def test_broadcast_then_device_resident_use(gpus):
    """BROADCAST runs the initialiser on every device; a following task on any device then finds its table initialised."""
    pool = ClDevicePool(ClDevicePoolType.DEVICE_COMPUTE_AT_WILL, SRC, False, 2)
    pool.addDevice(devices())
    table = ck.ClArray(N, np.int32)
    table.read, table.write = False, False             # device-resident: never uploaded or downloaded
    init = task(table, 1, "setTable", N, 256)
    init.type = ClTaskType.TASK_MESSAGE_BROADCAST | ClTaskType.TASK_MESSAGE_GLOBAL_SYNCHRONIZATION_LAST
    tp = ClTaskPool()
    tp.feed(init)
    outs = []
    for t in range(8):
        out = ck.ClArray(N, np.int32)
        out.writeOnly = True
        tp.feed(task(table.nextParam(out), 2, "useTable", N, 256))
        outs.append(out)
    pool.enqueueTaskPool(tp)
    pool.finish()
    for out in outs:
        assert np.array_equal(out.view, np.arange(N) + 7)
    assert sum(pool.tasksPerDevice()) == gpus + 8      # one broadcast copy per device
    pool.dispose()


@pytest.mark.parametrize("fine", [False, True], ids=["blocking", "fineGrained"])
def test_device_select_and_serial_mode_keep_a_run_on_one_device(gpus, fine):
    """A run of tasks works on device-resident data (upload in the first task, download in the last), which only holds
    together if all of its tasks land on the same device, in order.  SERIAL_MODE promises both in every pool.
    DEVICE_SELECT promises the device only: a fine-grained pool spreads the tasks of such a run over its queues, as the
    reference does (ClPipeline.cs:4946-4950 switches the round-robin off for serial mode alone), so ordered
    DEVICE_SELECT runs are checked in the blocking pool, where the single consumer of the device keeps them in order."""
    pool = ClDevicePool(ClDevicePoolType.DEVICE_COMPUTE_AT_WILL, SRC, fine, 3)
    pool.addDevice(devices())
    tp = ClTaskPool()
    results = []
    kinds = [(ClTaskType.TASK_MESSAGE_SERIAL_MODE_BEGIN, ClTaskType.TASK_MESSAGE_SERIAL_MODE_END)]
    if not fine:
        kinds.append((ClTaskType.TASK_MESSAGE_DEVICE_SELECT_BEGIN, ClTaskType.TASK_MESSAGE_DEVICE_SELECT_END))
    for run, (begin, end) in enumerate(kinds * 3):
        a = ck.ClArray(N, np.int32)
        a.view[:] = run
        for s in range(4):
            # the first task uploads, the middle ones touch only the device copy, the last one downloads; task() freezes
            # the array's flags as they are at this moment (ClArray.cs:1552-1583)
            a.read, a.write = (s == 0), (s == 3)
            tk = task(a, 50 + run, "addOne", N, 256)
            tk.type = (begin if s == 0 else 0) | (end if s == 3 else 0)
            tp.feed(tk)
        results.append(a)
    pool.enqueueTaskPool(tp)
    pool.finish()
    for run, a in enumerate(results):
        assert np.array_equal(a.view, np.full(N, run + 4, np.int32)), run
    pool.dispose()


def test_task_can_be_computed_directly(gpus):
    cr = ck.ClNumberCruncher(devices()[0], SRC)
    a = ck.ClArray(N, np.int32)
    a.view[:] = 5
    tk = task(a, 1, "addOne", N, 256)
    assert tk.compute(cr) == 0 and tk.compute(cr) == 0
    assert (a.view == 7).all()
