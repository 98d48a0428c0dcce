"""Pure-Python restatement of the reference's load balancer, independent of the C++ one in the product.

TEST INFRASTRUCTURE ONLY.  Follows /root/reference/Cekirdekler/Cekirdekler/HelperFunctions.cs:
  maxIndex                     :84-97
  performanceHistoryShiftOld   :119-133
  performanceHistoryAverage    :141-156
  loadBalance                  :190-280
and the first-call rule of Cores.compute (Cores.cs:569-596).  Python floats are IEEE doubles like C# `double`;
`int()` truncates toward zero like the C# `(int)` cast; `math.fmod`-style remainders are reproduced with
`_rem` because Python's % floors while C#'s truncates.
Parity pinning: the reference holds no golden vectors for the balancer — "parity unpinned";
tests/golden/balancer_cases.json records this restatement's outputs so later changes are caught.
"""
from __future__ import annotations

HISTORY_DEPTH = 10  # Cores.cs:1065


def _rem(a: int, b: int) -> int:
    """C#/C remainder (sign follows the dividend)."""
    r = abs(a) % abs(b)
    return -r if a < 0 else r


def _idiv(a: int, b: int) -> int:
    q = abs(a) // abs(b)
    return q if (a < 0) == (b < 0) else -q


def max_index(arr):
    result, value = 0, -100000000
    for i, v in enumerate(arr):
        if value < v:
            value, result = v, i
    return result


def new_history(devices: int):
    return [[0.0] * devices for _ in range(HISTORY_DEPTH)]


def load_balance(benchmark, smooth, history, total_range, global_ranges, step):
    """In-place on `global_ranges` and `history`; returns the raw throughputs (Cores.tmpThroughputs)."""
    n = len(global_ranges)
    tmp_ranges = [0] * n
    thr = [0.0] * n
    total_benchmark = 0.0
    for b in benchmark:
        total_benchmark += b
    total_benchmark += 0.01 * n
    total_throughput = 0.0
    for i in range(n):
        thr[i] = (total_benchmark / (benchmark[i] + 0.01)) * (global_ranges[i] + 1)
        total_throughput += thr[i]
    if total_throughput <= 0.0000001:
        total_throughput = 0.01
    normalized = [0.0] * n
    if smooth:
        for i in range(n):
            normalized[i] = thr[i] / total_throughput
        for k in range(HISTORY_DEPTH - 1):
            for j in range(n):
                history[k][j] = history[k + 1][j]
        for j in range(n):
            history[HISTORY_DEPTH - 1][j] = normalized[j]
        avg = [0.0] * n
        for k in range(HISTORY_DEPTH):
            for j in range(n):
                avg[j] += history[k][j]
        div = 1.0 / float(HISTORY_DEPTH)
        normalized = [a * div for a in avg]
    for i in range(n):
        share = normalized[i] if (smooth and history[0][0] > 0.00001) else (thr[i] / total_throughput)
        if global_ranges[i] != 0:
            tmp_ranges[i] = global_ranges[i] - int((global_ranges[i] - (total_range * share)) * 0.3)
        else:
            tmp_ranges[i] = int(total_range * (total_benchmark / (benchmark[i] + 0.01)) / total_throughput)
    for i in range(n):
        rem = _rem(tmp_ranges[i], step)
        if rem < _idiv(step, 2):
            global_ranges[i] = tmp_ranges[i] - rem
        else:
            global_ranges[i] = tmp_ranges[i] + (step - rem)
    while sum(global_ranges) > total_range:
        global_ranges[max_index(global_ranges)] -= step
    while sum(global_ranges) < total_range:
        global_ranges[max_index(global_ranges)] += step
    return thr


def load_balance_profile(benchmark, total_range, global_ranges, step, last_imbalance=0.0):
    """Independent restatement of the product's Profile mode (cekirdekler_b200/csrc/host/balance.cpp, profileBalance):
    no reference counterpart.  Device times are read as a piecewise-constant cost density over the range; the new
    boundaries cut its integral into equal parts; the move is damped.  In place on `global_ranges`; returns the imbalance
    (max/mean time of the active devices) the caller passes back as `last_imbalance` on the next call."""
    n = len(global_ranges)
    length, dens, total, mx = [], [], 0.0, 0.0
    for i in range(n):
        if global_ranges[i] > 0:
            t = benchmark[i] + 0.01
            length.append(float(global_ranges[i]))
            dens.append(t / float(global_ranges[i]))
            total += t
            mx = max(mx, t)
    if not length:
        return last_imbalance
    imbalance = mx * float(len(length)) / total
    target = total / float(n)
    want, seg, used, assigned = [0.0] * n, 0, 0.0, 0.0
    for d in range(n - 1):
        need, got = target, 0.0
        while seg < len(length):
            avail = (length[seg] - used) * dens[seg]
            if avail >= need - 1e-15:
                take = need / dens[seg] if dens[seg] > 0.0 else 0.0
                used += take
                got += take
                break
            need -= avail
            got += length[seg] - used
            seg += 1
            used = 0.0
        want[d] = got
        assigned += got
    want[n - 1] = float(total_range) - assigned
    gain = 0.8 if imbalance > 1.05 else 0.3
    if last_imbalance > 0.0 and imbalance > 1.02 * last_imbalance and imbalance > 1.05:
        gain = 0.5
    for i in range(n):
        tmp = int(float(global_ranges[i]) + gain * (want[i] - float(global_ranges[i])))
        rem = _rem(tmp, step)
        global_ranges[i] = max(0, tmp - rem if rem < _idiv(step, 2) else tmp + (step - rem))
    while sum(global_ranges) > total_range:
        global_ranges[max_index(global_ranges)] -= step
    while sum(global_ranges) < total_range:
        global_ranges[max_index(global_ranges)] += step
    return imbalance


def initial_split(total_range: int, devices: int):
    ranges = [_idiv(total_range, devices)] * devices
    s = sum(ranges)
    if s != total_range:
        ranges[0] += total_range - s
    return ranges


def balance_step(ranges, pipeline_enabled, blobs, local_range, total_range) -> int:
    b1 = all(r >= blobs * local_range for r in ranges)
    return blobs * local_range if (b1 and pipeline_enabled and total_range >= blobs * local_range) else local_range


class ComputeIdBalancer:
    """What Cores.compute does per call for one compute id (Cores.cs:544-613)."""

    def __init__(self, devices: int):
        self.devices = devices
        self.ranges = [0] * devices
        self.history = new_history(devices)

    def step(self, benchmark_ms, smooth, total_range, pipeline_enabled=False, blobs=4, local_range=256):
        first = all(r == 0 for r in self.ranges)
        if first:
            self.ranges = initial_split(total_range, self.devices)
            bench = [10.0] * self.devices
        else:
            bench = list(benchmark_ms)
        step = balance_step(self.ranges, pipeline_enabled, blobs, local_range, total_range)
        if _rem(total_range, step) != 0:
            step = local_range
        load_balance(bench, smooth, self.history, total_range, self.ranges, step)
        refs, acc = [], 0
        for r in self.ranges:
            refs.append(acc)
            acc += r
        return list(self.ranges), refs
