"""Balancer simulation on the real Mandelbrot row-cost profile (oracle, 1024 rows standing for 16384) — scratch tool."""
import ctypes as C
import sys

import numpy as np

sys.path.insert(0, "/root/repo")
from oracle import pyoracle as po
from cekirdekler_b200._native import host

W = H = 1024
view = po.mandelbrot_view(W, H, 4096)
it = po.mandelbrot(view, W * H).reshape(H, W).astype(np.float64)
row_cost = it.sum(axis=1) + 20.0 * W          # + per-pixel fixed cost
ITEMS = (16384 * 16384) // 4
STEP = 256
cum = np.concatenate([[0.0], np.cumsum(row_cost)])
cum *= 24.0 / cum[-1]                          # whole image = 24 ms on one device


def cost(a, b):
    """ms to compute items [a, b) on one device."""
    xa, xb = a / ITEMS * H, b / ITEMS * H
    return float(np.interp(xb, np.arange(H + 1), cum) - np.interp(xa, np.arange(H + 1), cum))


def times(ranges, rng, noise, xfer_ms_per_item=0.0, speeds=None):
    t, ref = [], 0
    for i, r in enumerate(ranges):
        c = cost(ref, ref + r) / (speeds[i] if speeds else 1.0)
        c = max(c, r * xfer_ms_per_item) + 0.02
        t.append(c * (1.0 + noise * rng.standard_normal()))
        ref += r
    return t


def profile_step(bench, ranges, total, step, state, gain_hi=1.0):
    """piecewise-constant cost density over the devices' segments → equal-cost boundaries"""
    n = len(ranges)
    b = [x + 0.01 for x in bench]
    act = [i for i in range(n) if ranges[i] > 0]
    C_tot = sum(b[i] for i in act)
    mx = max(b[i] for i in act)
    imb = mx * len(act) / C_tot
    target = C_tot / n
    # boundaries
    new = [0.0] * n
    seg, pos_in_seg, acc_dev = 0, 0.0, 0.0      # walking pointer: segment index, items consumed in it
    segs = [(ranges[i], b[i] / ranges[i]) for i in act]
    k = 0
    out = []
    cur_len = 0.0
    need = target
    si, used = 0, 0.0
    for dev in range(n - 1):
        need = target
        length = 0.0
        while si < len(segs):
            L, d = segs[si]
            avail = (L - used) * d
            if avail >= need - 1e-15:
                take = need / d if d > 0 else 0.0
                used += take
                length += take
                break
            need -= avail
            length += L - used
            si += 1
            used = 0.0
        out.append(length)
    out.append(total - sum(out))
    gain = gain_hi if imb > 1.05 else 0.3
    if state.get("last") and imb > 1.02 * state["last"] and imb > 1.05:
        gain = min(gain, 0.5)
    state["last"] = imb
    tmp = [int(ranges[i] + gain * (out[i] - ranges[i])) for i in range(n)]
    res = []
    for v in tmp:
        rem = v % step
        res.append(v - rem if rem < step // 2 else v + step - rem)
    while sum(res) > total:
        res[int(np.argmax(res))] -= step
    while sum(res) < total:
        res[int(np.argmax(res))] += step
    return res, imb


def run(mode, D, calls=25, noise=0.01, seed=0, xfer=0.0, speeds=None, gain_hi=1.0):
    rng = np.random.default_rng(seed)
    h = host()
    imbs = []
    if mode in (0, 1):
        b = h.ckh_balancer_create(D, mode)
        ranges = (C.c_int32 * D)(*([0] * D))
        t = (C.c_double * D)(*([10.0] * D))
        for c in range(calls):
            h.ckh_balancer_step(b, t, 1, ITEMS, ranges, 0, 4, STEP)
            tt = times(list(ranges), rng, noise, xfer, speeds)
            imbs.append(max(tt) / (sum(tt) / D))
            for i in range(D):
                t[i] = tt[i]
        h.ckh_balancer_destroy(b)
    else:
        ranges = [ITEMS // D // STEP * STEP] * D
        ranges[0] += ITEMS - sum(ranges)
        state = {}
        for c in range(calls):
            tt = times(ranges, rng, noise, xfer, speeds)
            imbs.append(max(tt) / (sum(tt) / D))
            ranges, _ = profile_step(tt, ranges, ITEMS, STEP, state, gain_hi)
    return imbs


def first_stable(imbs, tol=1.05, hold=3):
    for k in range(len(imbs)):
        if all(v <= tol for v in imbs[k:k + hold + 1]) and k + hold < len(imbs):
            return k
    return None


if __name__ == "__main__":
    for D in (2, 4, 8):
        for label, kw in (("compute-bound", {}), ("with host link", {"xfer": 19.5 / ITEMS}), ("one slow device", {"speeds": [1.0] * (D - 1) + [0.6]})):
            for mode, name in ((0, "reference"), (1, "adaptive"), (2, "profile g1.0"), (3, "profile g0.8")):
                ks, finals = [], []
                for seed in range(5):
                    imbs = run(mode if mode < 2 else 2, D, seed=seed, gain_hi=1.0 if mode == 2 else 0.8, **kw)
                    ks.append(first_stable(imbs))
                    finals.append(np.mean(imbs[-5:]))
                print(f"D={D} {label:16s} {name:13s} converge at {ks}  final imbalance {np.mean(finals):.3f}")

def extra():
    global cum
    for D, label, kw in ((2, "slow 0.3", {"speeds": [1.0, 0.3]}), (8, "two slow", {"speeds": [0.5, 1, 1, 1, 1, 1, 1, 0.3]}), (8, "noise 3%", {"noise": 0.03}),
                         (3, "compute", {}), (5, "slow mid", {"speeds": [1, 1, 0.4, 1, 1]}), (8, "noise 3% link", {"noise": 0.03, "xfer": 19.5 / ITEMS})):
        for mode, name, g in ((0, "reference", 0), (1, "adaptive", 0), (2, "profile g0.8", 0.8), (2, "profile g0.7", 0.7)):
            ks, finals = [], []
            for seed in range(5):
                imbs = run(mode, D, seed=seed, gain_hi=g, **kw)
                ks.append(first_stable(imbs)); finals.append(np.mean(imbs[-5:]))
            print(f"D={D} {label:14s} {name:13s} converge at {ks}  final imbalance {np.mean(finals):.3f}")
    # uniform workload (n-body like)
    save = cum.copy()
    cum = np.linspace(0, 24.0, H + 1)
    for D, label, kw in ((8, "uniform", {}), (8, "uniform slow", {"speeds": [1, 1, 1, 0.5, 1, 1, 1, 1]}), (2, "uniform slow", {"speeds": [1, 0.25]})):
        for mode, name, g in ((0, "reference", 0), (1, "adaptive", 0), (2, "profile g0.8", 0.8)):
            ks, finals = [], []
            for seed in range(5):
                imbs = run(mode, D, seed=seed, gain_hi=g, **kw)
                ks.append(first_stable(imbs)); finals.append(np.mean(imbs[-5:]))
            print(f"D={D} {label:14s} {name:13s} converge at {ks}  final imbalance {np.mean(finals):.3f}")
    cum = save

if __name__ == "__main__":
    print("---- extra")
    extra()
