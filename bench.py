#!/usr/bin/env python
"""bench.py — the hot path's headline benchmark on B200.

    python bench.py --gpus N --steps K --warmup W              (N>1: launched by torchrun, one rank per GPU)
    python bench.py --impl reference --gpus N --steps K --warmup W   (the CPU path on the box's host cores)

Headline workload (BASELINE.json configs[2], the configuration the Mpix/s metric is quoted on; it fits one GPU):
Mandelbrot 16384 x 16384, 4096 max iterations, view [-2,1] x [-1.5,1.5], through ClNumberCruncher.compute().
A "step" is one compute() over the whole image.  Three measurements per run:
  value      Mpix/s with everything resident in HBM (the output stays in the device buffer): CUDA-event time of the
             device span of each compute(), summed over the K steps, max over ranks.
  e2e        the same metric through the public API with HOST buffers: every step uploads the view array and downloads
             the iteration counts into pinned host memory (event-driven pipelining, 16 blobs, hides most of the download).
  roofline   achieved FP32 rate of the Mandelbrot kernel = 8 FLOP x (sum of iteration counts) / kernel time, against
             2 x 128 lanes x SMs x max SM clock (no tensor cores on this path; the bound is the FP32 FMA pipe).
plus the other BASELINE workloads (streaming SAXPY / vector-add 2^28 floats, n-body 128 Ki bodies) under "workloads",
each with its own roofline entry, and the CPU oracle timed on a bounded sample ("cpu_baseline").
Multi-GPU: under torchrun every rank drives one GPU as one device of the split (cekirdekler_b200/spmd.py); the ranks
exchange only their device times, and the balancer's iterations-to-converge is reported.  The total work is fixed, so
scaling is "strong".
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

_REAL_STDOUT = None


def emit(obj):
    """The one JSON line, on the process's original stdout."""
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(obj) + "\n")
    out.flush()


METRIC = "mandelbrot_mpix_per_s"
UNIT = "Mpix/s"
W = H = 16384
MAX_ITER = 4096
WORKLOAD = f"mandelbrot {W}x{H} maxIter {MAX_ITER} view [-2,1]x[-1.5,1.5] (BASELINE configs[2])"   # the same string in both arms
E = 1  # pixels per work-item: 1 for `mandelbrot1` (the default), 4 for `mandelbrot`; set from --mandel-kernel in main()
MK = "mandelbrot1"
LOCAL = int(os.environ.get("CK_MANDEL_LOCAL", "256"))   # work-group size of the headline kernel (experiment hook; one B200: 64 -> 84.7 %, 128 and 256 -> 86.7 %, 512 -> 85.8 %, 1024 -> 84.8 % of FP32)


def peaks():
    p = {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0, "source": "fallback (B200_PROFILING.md)"}
    f = ROOT / "MEASURED_PEAKS.json"
    if f.exists():
        try:
            j = json.loads(f.read_text())
            p.update(hbm_gbs=float(j["hbm_gbs"]), sm_max_mhz=float(j["sm_max_mhz"]), source="MEASURED_PEAKS.json")
        except Exception:
            pass
    return p


class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int):
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {
            nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
            nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.05)

    def __enter__(self):
        if self.nv:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._t:
            self._t.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def ncu_traffic(kernel, summary="r1_ncu_summary.json"):
    """DRAM bytes per launch of `kernel` from the committed `ncu --set full` capture (profiles/<summary>), or None."""
    f = ROOT / "profiles" / summary
    try:
        return float(json.loads(f.read_text())[kernel]["traffic_B"])
    except Exception:
        return None


def dist_env():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


# ------------------------------------------------------------------------------------------------ CPU arm
def cpu_sample(po, view, seconds_target=12.0):
    """Times the oracle (all host threads) on a strided sample of the C3 image sized for ~seconds_target.
    The stride is odd, hence coprime with the 2^14-pixel row length: the sample walks over every column and row band
    and sees the image's cost distribution (a probe with the same property calibrates the sample size)."""
    npix = W * H
    probe_stride, probe = 4099, 1 << 15
    t0 = time.perf_counter()
    po.mandelbrot_sample(view, 777, probe_stride, probe)
    pix_per_s = probe / max(time.perf_counter() - t0, 1e-3)
    count = int(max(probe, min(npix // 3, pix_per_s * seconds_target)))
    stride = max(3, (npix - 1000) // count) | 1
    count = min(count, (npix - 1000) // stride)   # the last sample index stays inside the image
    buf = np.zeros(count, np.int32)
    t0 = time.perf_counter()
    iters = po.mandelbrot_sample(view, 777, stride, count, buf)
    dt = time.perf_counter() - t0
    return {"pixels": count, "seconds": dt, "iters": int(iters), "stride": stride, "offset": 777, "values": buf}


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return
    from oracle import pyoracle as po

    po.use_all_cores()
    view = po.mandelbrot_view(W, H, MAX_ITER)
    per_step = max(2.0, min(20.0, 90.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        cpu_sample(po, view, per_step / 4)
    tot_pix, tot_s = 0, 0.0
    for _ in range(args.steps):
        s = cpu_sample(po, view, per_step)
        tot_pix += s["pixels"]
        tot_s += s["seconds"]
    v = tot_pix / tot_s / 1e6
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": tot_s / args.steps * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD,
                   "note": "the reference's PoCL/C# path cannot run on this box (no CLR, no OpenCL); this is the CPU "
                           "restatement oracle/oracle.cpp on all host threads"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": po.threads(), "kind": "port",
                         "sample": f"{tot_pix // args.steps} pixels per step, an odd-strided sample across the whole {W}x{H} image"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ------------------------------------------------------------------------------------------------ GPU arm helpers
def first_stable(imb, tol=1.05, hold=3):
    for k in range(len(imb) - hold):
        if all(v <= tol for v in imb[k:k + hold + 1]):
            return k + 1
    return None


def bench_hostlink(L, ordinals=(0,), mib=512, reps=4, barrier=None):
    """Host-link roofline for the streamed workloads: pinned H2D, D2H and both at once, through the C ABI's own transfers, on
    every device of `ordinals` AT THE SAME TIME (one pinned array pair, one queue pair per device).  With one process per
    GPU every rank calls this with its own device and `barrier`, so that all ranks copy together; the caller sums the ranks.
    Returns this process's aggregate GB/s and the per-device figures."""
    n = mib * (1 << 20) // 4
    lst = L.platformList()
    plat = L.createPlatform(lst, 0)
    L.deletePlatformList(lst)
    devs = []
    for o in ordinals:
        dev = L.createDevice(plat, L.CODE_GPU(), o)
        ctx = L.createContext(plat, dev)
        q1, q2 = L.createCommandQueue(ctx, dev, 0), L.createCommandQueue(ctx, dev, 0)
        a1, a2 = L.createArray(n, 4096, 0), L.createArray(n, 4096, 0)
        h1, h2 = L.alignedArrHead(a1), L.alignedArrHead(a2)
        b1, b2 = L.createBuffer(ctx, n, 0, 0, h1, 1, 0, 0), L.createBuffer(ctx, n, 0, 0, h2, 1, 0, 0)
        devs.append(dict(dev=dev, ctx=ctx, q1=q1, q2=q2, a1=a1, a2=a2, h1=h1, h2=h2, b1=b1, b2=b2))
    sync = barrier or (lambda: None)

    def drain():
        for d in devs:
            L.finish(d["q1"]), L.finish(d["q2"])

    def timed(fn):
        """Seconds per repetition for the whole set, and per device (each device's own queues drained in turn)."""
        for d in devs:
            fn(d)
        drain()
        sync()
        t0 = time.perf_counter()
        for _ in range(reps):
            for d in devs:
                fn(d)
        per = []
        for d in devs:
            L.finish(d["q1"]), L.finish(d["q2"])
            per.append((time.perf_counter() - t0) / reps)
        sync()
        return (time.perf_counter() - t0) / reps, per

    byts = n * 4
    out = {"devices": list(ordinals), "MiB_per_copy": mib}
    for key, fn in (("h2d", lambda d: L.writeToBuffer(d["q1"], d["b1"], d["h1"])),
                    ("d2h", lambda d: L.readFromBuffer(d["q2"], d["b2"], d["h2"])),
                    ("bidirectional_each", lambda d: (L.writeToBuffer(d["q1"], d["b1"], d["h1"]), L.readFromBuffer(d["q2"], d["b2"], d["h2"])))):
        t, per = timed(fn)
        out[key + "_GBps"] = byts * len(devs) / t / 1e9          # aggregate over this process's devices
        if len(devs) > 1:
            out[key + "_GBps_per_device"] = [round(byts / p / 1e9, 2) for p in per]
    for d in devs:
        for h, f in ((d["b1"], L.deleteBuffer), (d["b2"], L.deleteBuffer), (d["a1"], L.deleteArray), (d["a2"], L.deleteArray),
                     (d["q1"], L.deleteCommandQueue), (d["q2"], L.deleteCommandQueue), (d["ctx"], L.deleteContext), (d["dev"], L.deleteDevice)):
            f(h)
    L.deletePlatform(plat)
    return out


def bench_pipeline(ck, kernels, po, n_gpus, frames=24):
    """C5: blur -> Sobel -> histogram on 3840x2160 float frames, one stage per GPU (all on GPU 0 when fewer), NVLink
    hand-off vs the reference's host bounce; plus the same frames as independent tasks through the greedy device pool."""
    from cekirdekler_b200.pipeline import ClPipelineStage
    from cekirdekler_b200.pool import ClDevicePool, ClDevicePoolType, ClTaskPool, task

    Wf, Hf = 3840, 2160
    n = Wf * Hf
    devs = ck.ClPlatforms.all().gpus()
    d = lambda i: devs[i % max(1, n_gpus)]
    RB = 30
    src = kernels.image_pipeline_source()
    gs, ls = kernels.strip_launch(Wf, Hf, RB)
    frame = ck.ClFloatArray(n)
    frame.view[:] = po.uniform(0xF00D, n, 0, 1)
    dims = np.array([Wf, Hf, RB], np.int32)
    dims2 = np.array([Wf, Hf], np.int32)
    b, s, hexp = np.zeros(n, np.float32), np.zeros(n, np.float32), np.zeros(256, np.uint32)
    po.blur5x5(frame.view, b, dims2)
    po.sobel(b, s, dims2)
    po.histogram(s, hexp)
    out = {"frame": f"{Wf}x{Hf} float", "stages_on": [d(i).ordinals[0] for i in range(3)]}
    for host_fwd in (False, True):
        blur, sob, hist = ClPipelineStage(), ClPipelineStage(), ClPipelineStage()
        blur.addDevices(d(0)); blur.addKernels(src, "blurStage", [gs], [ls])
        blur.addInputBuffers((np.float32, n), (np.int32, 3)); blur.addOutputBuffers((np.float32, n), (np.int32, 3))
        sob.addDevices(d(1)); sob.addKernels(src, "sobelStage", [gs], [ls])
        sob.addInputBuffers((np.float32, n), (np.int32, 3)); sob.addOutputBuffers((np.float32, n), (np.int32, 3)); sob.appendToStage(blur)
        hist.addDevices(d(2)); hist.addKernels(src, "histStage", [592 * 256], [256])
        hist.addInputBuffers((np.float32, n), (np.int32, 3)); hist.addHiddenBuffers((np.uint32, 2049))
        hist.addOutputBuffers((np.uint32, 256)); hist.appendToStage(sob)
        pipe = hist.makePipeline(hostForwarding=host_fwd)
        assert pipe.errorCode() == 0, pipe.errorMessage()
        res = np.zeros(256, np.uint32)
        for _ in range(8):                                   # fill the pipeline (latency 2S+1 pushes)
            pipe.pushData([frame.view, dims], [res])
        ok = bool(np.array_equal(res, hexp))
        t0 = time.perf_counter()
        for _ in range(frames):
            pipe.pushData([frame.view, dims], [res])         # mode B: every frame comes from (pinned) host memory
        dt = (time.perf_counter() - t0) / frames
        one_device = len(set(out["stages_on"])) == 1       # then the "hand-off" is a copy inside one GPU's HBM, not NVLink
        key = "host_bounce" if host_fwd else ("device_copy_same_gpu" if one_device else "nvlink_p2p")
        out[key] = {"frames_per_s": 1.0 / dt, "ms_per_frame": dt * 1e3, "histogram_exact": ok and bool(np.array_equal(res, hexp))}
        if not host_fwd:
            t0 = time.perf_counter()
            for _ in range(frames):
                pipe.pushData()                              # mode A: frames already resident on the first device
            dt = (time.perf_counter() - t0) / frames
            out[key + "_resident"] = {"frames_per_s": 1.0 / dt, "ms_per_frame": dt * 1e3,
                                          "GBps_algorithmic_hbm": 165.9e6 / dt / 1e9}
        pipe.dispose()
    # task pool: independent frames (blur only), greedy over the devices
    from cekirdekler_b200.pool import ClTaskType

    pool = ClDevicePool(ClDevicePoolType.DEVICE_COMPUTE_AT_WILL, kernels.source("image"), True, 3)
    pool.addDevice(devs[0:max(1, n_gpus)])
    dims_a = ck.ClIntArray(3)
    dims_a.view[:] = dims
    dims_a.readOnly = True
    frame.readOnly = True
    outs, tasks = [], []
    ntasks = 32
    for t in range(ntasks):
        o = ck.ClFloatArray(n)
        o.read, o.write, o.writeAll = False, True, True          # the strip kernel's work-items are not pixels: whole frame back
        outs.append(o)
        tasks.append(task(frame.nextParam(o, dims_a), 1 + t % 3, "blurStrips", gs, ls))
    # Warm pass: every task on EVERY device (broadcast), so that each device already owns the buffers of whichever tasks the
    # greedy split hands it later — a buffer made inside the timed pass is a cudaMalloc, which drains the whole GPU.
    warm = ClTaskPool()
    for tk in tasks:
        tk.type = ClTaskType.TASK_MESSAGE_BROADCAST
        warm.feed(tk)
    pool.enqueueTaskPool(warm)
    pool.finish()
    tp = ClTaskPool()
    for tk in tasks:
        tk.type = ClTaskType.TASK_MESSAGE_DEFAULT
        tp.feed(tk)
    pool.enqueueTaskPool(tp)
    pool.finish()                                            # one untimed greedy pass
    pool.enqueueTaskPool(tp)
    ms = pool.finish()
    out["task_pool"] = {"tasks": ntasks, "kernel": "blurStrips", "ms_total": ms, "frames_per_s": ntasks / (ms * 1e-3),
                        "tasks_per_device": pool.tasksPerDevice(), "exact": bool(np.array_equal(outs[-1].view, b)),
                        "host_link_GBps": 2.0 * 4 * n * ntasks / (ms * 1e-3) / 1e9,
                        "note": "each task uploads a 33 MB frame and downloads a 33 MB result (host-link bound); buffers of every "
                                "task exist on every device before the timed pass"}
    pool.dispose()
    return out


def bench_image_kernels(ck, kernels, po, dev, pk, sets=6, rounds=3):
    """K5a/b/c alone and fused against the HBM roofline (8 B/px, 8 B/px, 4 B/px + 1 KiB; fused 4 B/px + 1 KiB).
    Every array stays on the device (flags cleared after the first call); `sets` independent frame sets (6 x 3 x 33 MB =
    597 MB >> 126 MB L2) are cycled so that no launch finds its frame in L2.  Two times per kernel: the CUDA-event span of
    one launch (what the balancer sees; includes the launch's ramp and tail on an otherwise idle GPU), and the time per
    launch when the launches follow each other without a host wait (enqueue mode), which is how a pipeline runs them."""
    Wf, Hf, RB = 3840, 2160, 30
    n = Wf * Hf
    cr = ck.ClNumberCruncher(dev, kernels.source("image"))
    assert cr.errorCode() == 0, cr.errorMessage()
    dims = ck.ClIntArray(3)
    dims.view[:] = [Wf, Hf, RB]
    dims.readOnly = True
    frame = po.uniform(0xF00D, n, 0, 1)
    b_exp, s_exp, h_exp = np.zeros(n, np.float32), np.zeros(n, np.float32), np.zeros(256, np.uint32)
    dims_h = np.array([Wf, Hf], np.int32)
    po.blur5x5(frame, b_exp, dims_h)
    po.sobel(b_exp, s_exp, dims_h)
    po.histogram(s_exp, h_exp)
    S = []
    for _ in range(sets):
        a, b, c, h, acc = ck.ClFloatArray(n), ck.ClFloatArray(n), ck.ClFloatArray(n), ck.ClUIntArray(256), ck.ClUIntArray(2049)
        a.view[:] = frame
        S.append((a, b, c, h, acc))
    gs, ls = kernels.strip_launch(Wf, Hf, RB)
    hg = int(os.environ.get("CK_HIST_GROUPS", "592"))       # work-groups of histogramStrips (experiment hook)
    # name: (kernel names, argument group, compute id, global range, local, elements per item of each argument, algorithmic bytes)
    calls = {"blurStrips": ("blurStrips", lambda s: s[0].nextParam(s[1], dims), 11, gs, ls, [1, 1, 1], 8.0 * n),
             "sobelStrips": ("sobelStrips", lambda s: s[1].nextParam(s[2], dims), 12, gs, ls, [1, 1, 1], 8.0 * n),
             "histogramStrips": ("histogramStrips", lambda s: s[2].nextParam(s[3], dims, s[4]), 13, hg * 256, 256, [1, 1, 1, 1], 4.0 * n + 1024),
             "blurSobelHistogram": ("blurSobelHistogram", lambda s: s[0].nextParam(s[3], dims, s[4]), 14, gs, ls, [1, 1, 1, 1], 4.0 * n + 1024),
             "gaussianBlur": ("gaussianBlur", lambda s: s[0].nextParam(s[1], dims), 1, n, 256, [1, 1, 1], 8.0 * n),
             "gaussianBlur8": ("gaussianBlur8", lambda s: s[0].nextParam(s[1], dims), 4, n // 8, 256, [1, 8, 1], 8.0 * n),
             "sobel": ("sobel", lambda s: s[1].nextParam(s[2], dims), 2, n, 256, [1, 1, 1], 8.0 * n),
             "histogram": ("histogramZero histogram", lambda s: s[2].nextParam(s[3]), 3, n // 16, 256, [16, 1], 4.0 * n + 1024)}
    traffic = {k: ncu_traffic(k, "r2_ncu_summary.json") for k in calls}
    out = {"rows_per_band": RB}
    # first pass: upload the source frame of every set once and run the chain, so that every later launch finds real data on the device
    for s in S:
        assert s[0].nextParam(s[1], dims).compute(cr, 11, "blurStrips", gs, ls, readWritesReady=[" read ", "", " read "]) == 0
        assert s[1].nextParam(s[2], dims).compute(cr, 12, "sobelStrips", gs, ls, readWritesReady=["", "", ""]) == 0
    for k, (names, group, cid, rng_items, local, epi, nbytes) in calls.items():
        rw = [""] * len(epi)
        ts = []
        for r in range(rounds + 1):
            for s in S:
                assert group(s).compute(cr, cid, names, rng_items, local, readWritesReady=rw, elementsPerItemReady=epi) == 0
                if r > 0:
                    ts.append(cr.benchmarks(cid)[0] * 1e-3)
        dt = float(np.mean(ts))
        cr.enqueueMode = True                     # launches follow each other; the host waits once at the end
        t0 = time.perf_counter()
        for r in range(rounds * 3):
            for s in S:
                assert group(s).compute(cr, cid, names, rng_items, local, readWritesReady=rw, elementsPerItemReady=epi) == 0
        cr.enqueueMode = False
        cr.sync()
        dq = (time.perf_counter() - t0) / (rounds * 3 * len(S))
        out[k] = {"us_per_frame": dt * 1e6, "us_per_frame_back_to_back": dq * 1e6, "GBps": nbytes / dt / 1e9,
                  "roofline": {"bound": "hbm", "achieved": nbytes / dt / 1e9, "peak": pk["hbm_gbs"], "unit": "GB/s",
                               "frac": nbytes / dt / 1e9 / pk["hbm_gbs"], "frac_back_to_back": nbytes / dq / 1e9 / pk["hbm_gbs"],
                               "traffic": traffic[k], "algorithmic_bytes": nbytes}}
    # parity of what is left on the device after the timed launches (set 0): downloads only
    s = S[0]
    for arr in (s[1], s[2], s[3]):
        arr.view[:] = 0
    assert s[0].nextParam(s[1], dims).compute(cr, 11, "blurStrips", gs, ls, readWritesReady=["", " write all ", ""]) == 0
    assert s[1].nextParam(s[2], dims).compute(cr, 12, "sobelStrips", gs, ls, readWritesReady=["", " write all ", ""]) == 0
    assert s[0].nextParam(s[3], dims, s[4]).compute(cr, 14, "blurSobelHistogram", gs, ls, readWritesReady=["", " write all ", "", ""]) == 0
    fused_ok = bool(np.array_equal(s[3].view, h_exp))
    s[3].view[:] = 0
    assert s[2].nextParam(s[3], dims, s[4]).compute(cr, 13, "histogramStrips", 592 * 256, 256, readWritesReady=["", " write all ", "", ""]) == 0
    out["parity"] = {"frame": f"{Wf}x{Hf}", "against": "oracle_blur5x5 / oracle_sobel / oracle_histogram, every pixel and bin",
                     "blurStrips_mismatches": int((s[1].view.view(np.uint32) != b_exp.view(np.uint32)).sum()),
                     "sobelStrips_mismatches": int((s[2].view.view(np.uint32) != s_exp.view(np.uint32)).sum()),
                     "histogramStrips_exact": bool(np.array_equal(s[3].view, h_exp)), "blurSobelHistogram_exact": fused_ok}
    for s in S:
        for x in s:
            x.dispose()
    dims.dispose()
    cr.dispose()
    return out


def bench_c1(ck, kernels, po, dev, pk, sms, reps=200):
    """C1 (BASELINE.json configs[0]): Mandelbrot 1024 x 1024, 256 iterations, one compute id, one GPU — the size at which the
    per-call cost of compute() matters (SURVEY.md §7 hard part 6: a ~50 us kernel).  Reported per mode: wall time per
    compute() call, the device span of the call (CUDA events: upload .. download), and their difference = what the host
    side of one call costs; plus the kernel alone with everything resident."""
    Wc = Hc = 1024
    npx = Wc * Hc
    cr = ck.ClNumberCruncher(dev, kernels.source("mandelbrot"))
    assert cr.errorCode() == 0, cr.errorMessage()
    view, out = ck.ClArray(8, np.float32), ck.ClArray(npx, np.int32)
    view.view[:] = po.mandelbrot_view(Wc, Hc, 256)
    view.readOnly = True
    out.writeOnly = True
    out.numberOfElementsPerWorkItem = E
    g = view.nextParam(out)
    exp = po.mandelbrot(view.view, npx)
    res = {"pixels": npx, "max_iter": 256, "iterations_total": int(exp.sum(dtype=np.int64))}

    def run(label, call, cid):
        for _ in range(10):
            assert call() == 0
        wall, span = [], []
        for _ in range(reps):
            t0 = time.perf_counter()
            assert call() == 0
            wall.append(time.perf_counter() - t0)
            span.append(cr.benchmarks(cid)[0] * 1e-3)
        w, d = float(np.median(wall)), float(np.median(span))
        res[label] = {"us_per_call_wall": w * 1e6, "us_per_call_device_span": d * 1e6, "us_host_overhead_per_call": (w - d) * 1e6,
                      "Mpix_per_s": npx / w / 1e6}

    out.view[:] = -1
    run("plain", lambda: g.compute(cr, 1, MK, npx // E, 256), 1)
    ok_plain = bool(np.array_equal(out.view, exp))
    out.view[:] = -1
    run("event_pipeline_4_blobs", lambda: g.compute(cr, 2, MK, npx // E, 256, 0, True, ck.PIPELINE_EVENT, 4), 2)
    ok_pipe = bool(np.array_equal(out.view, exp))
    del res["event_pipeline_4_blobs"]["us_per_call_device_span"], res["event_pipeline_4_blobs"]["us_host_overhead_per_call"]  # host clock there
    rw = ["", ""]
    run("resident_kernel_only", lambda: g.compute(cr, 3, MK, npx // E, 256, 0, False, True, 4, readWritesReady=rw,
                                                  elementsPerItemReady=[1, E]), 3)
    # the host side alone: enqueue mode (no wait per call) on one work-group, so that the GPU is never what is waited for
    cr.enqueueMode = True
    tiny = lambda: g.compute(cr, 4, MK, 256, 256, 0, False, True, 4, readWritesReady=rw, elementsPerItemReady=[1, E])
    for _ in range(200):
        assert tiny() == 0
    cr.enqueueMode = False
    cr.enqueueMode = True
    t0 = time.perf_counter()
    for _ in range(5000):
        tiny()
    dt_host = (time.perf_counter() - t0) / 5000
    cr.enqueueMode = False
    res["host_us_per_call_enqueue_mode"] = {"us": dt_host * 1e6, "note": "compute() through the Python binding and the C++ layer down to "
                                            "cudaLaunchKernel, no wait: 5000 calls of one work-group in enqueue mode"}
    k = res["resident_kernel_only"]["us_per_call_device_span"] * 1e-6
    peak = 2.0 * 128 * sms * pk["sm_max_mhz"] * 1e6 / 1e12
    ach = 8.0 * res["iterations_total"] / k / 1e12
    res["resident_kernel_only"]["roofline"] = {"bound": "fp32", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                                               "note": "a 1024^2 x 256 launch is ~4096 work-groups of very unequal length: launch ramp and tail included"}
    res["parity"] = {"pixels_checked": npx, "mismatches_plain": int((out.view != exp).sum()) if not ok_pipe else 0,
                     "exact_plain": ok_plain, "exact_event_pipeline": ok_pipe}
    view.dispose(); out.dispose(); cr.dispose()
    return res


def bench_saxpy(ck, kernels, dev, shim, pk, steps, warmup):
    """C2: streaming SAXPY, 2^28 floats, 16 blobs: resident (HBM roofline) and streamed (host-link roofline)."""
    n = 1 << 28
    cr = ck.ClNumberCruncher(dev, kernels.source("stream"))
    assert cr.errorCode() == 0, cr.errorMessage()
    alpha, x, y = ck.ClFloatArray(1), ck.ClFloatArray(n), ck.ClFloatArray(n)
    alpha[0] = 2.0
    x.view[:] = 3.0
    y.view[:] = (np.arange(n, dtype=np.int64) % (1 << 20)).astype(np.float32)
    alpha.readOnly = True
    for a in (x, y):
        a.partialRead, a.read = True, False
        a.numberOfElementsPerWorkItem = 4
    x.readOnly = True
    out = {}
    g = alpha.nextParam(x, y)
    # streamed: both pipeline modes, host buffers every call
    for mode, ptype in (("event", ck.PIPELINE_EVENT), ("driver", ck.PIPELINE_DRIVER)):
        ts = []
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            assert g.compute(cr, 10 + int(ptype), "saxpy4", n // 4, 256, 0, True, ptype, 16) == 0
            if i >= warmup:
                ts.append(time.perf_counter() - t0)
        dt = float(np.mean(ts))
        out[f"streamed_{mode}"] = {"ms_per_step": dt * 1e3, "GBps_algorithmic": 12.0 * n / dt / 1e9,
                                   "h2d_GBps": 8.0 * n / dt / 1e9, "d2h_GBps": 4.0 * n / dt / 1e9}
    calls = 2 * (warmup + steps)
    # every element of every blob of both lanes: y = (i mod 2^20) + 6 per call, small integers, so the sum is exact in f32
    bad_blobs, mismatches, blob = [], 0, n // 16
    base = (np.arange(blob, dtype=np.int64) % (1 << 20)).astype(np.float32) + np.float32(6.0 * calls)   # blob is a multiple of 2^20
    for k in range(16):
        d = int(np.count_nonzero(y.view[k * blob:(k + 1) * blob] != base))
        mismatches += d
        if d:
            bad_blobs.append(k)
    out["parity"] = {"elements_checked": n, "blobs_checked": 16, "mismatches": mismatches, "blobs_with_mismatches": bad_blobs,
                     "against": "y0 + 2*3*calls, exact in f32 (what oracle_saxpy gives on these integers)"}
    # resident: flags cleared after the first call (as Kamera.cs:271-272 does) → no transfers, kernel only
    g_res = alpha.nextParam(x, y)
    rw = ["", "", ""]
    ts = []
    for i in range(max(3, warmup) + steps):
        assert g_res.compute(cr, 20, "saxpy4", n // 4, 256, 0, False, True, 4, readWritesReady=rw, elementsPerItemReady=[1, 4, 4]) == 0
        if i >= max(3, warmup):
            ts.append(cr.benchmarks(20)[0] * 1e-3)
    dt = float(np.mean(ts))
    gbs = 12.0 * n / dt / 1e9
    out["resident"] = {"ms_per_step": dt * 1e3, "GBps": gbs,
                       "roofline": {"bound": "hbm", "achieved": gbs, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": gbs / pk["hbm_gbs"],
                                    "traffic": ncu_traffic("saxpy"), "algorithmic_bytes": 12.0 * n,
                                    "peak_source": pk["source"] + " (a 1:1 copy; SAXPY's 2 reads : 1 write moves more)"}}
    for a in (alpha, x, y):
        a.dispose()
    cr.dispose()
    return out


def own_bodies(cr, cid, per_body, spmd_ctx):
    """(first body, bodies) of every share this process computes."""
    r, f = cr.globalRanges(cid), cr.globalReferences(cid)
    idx = [spmd_ctx[0]] if spmd_ctx is not None else range(len(r))
    return [(f[i] // per_body, r[i] // per_body) for i in idx if r[i] > 0]


def nbody_parity(po, pos_before, vel_before, pos_after, vel_after, n, slices, total=1024, block=64):
    """One oracle step on `total` bodies spread over `slices`, each against all n bodies; worst relative position error."""
    params = np.array([1e-3, 1e-4, n, 0], np.float32)
    worst, worst_v, checked = 0.0, 0.0, 0
    per_slice = max(1, total // block // max(1, len(slices)))
    after, vafter = pos_after.reshape(n, 4), vel_after.reshape(n, 4)
    for (first, count) in slices:
        for k in range(per_slice):
            if count < block:
                continue
            f = first + (k * (count - block)) // max(1, per_slice - 1)
            v, o = vel_before.copy(), np.zeros(4 * n, np.float32)
            po.nbody3d_step(pos_before, v, o, params, f, block)
            exp = o.reshape(n, 4)[f:f + block]
            rel = np.abs(after[f:f + block, :3] - exp[:, :3]) / np.maximum(np.abs(exp[:, :3]), 1e-3)
            worst = max(worst, float(rel.max()))
            vexp = v.reshape(n, 4)[f:f + block, :3]
            worst_v = max(worst_v, float(np.abs(vafter[f:f + block, :3] - vexp).max() / max(float(np.abs(vexp).max()), 1e-12)))
            checked += block
    return {"bodies_checked": checked, "against": f"oracle_nbody3d_step, each body against all {n}", "max_rel_err_positions": worst,
            "max_rel_err_velocities": worst_v, "tolerance": 1e-4, "ok": bool(checked > 0 and worst <= 1e-4)}


def bench_nbody(ck, kernels, po, devs, pk, sms, steps, warmup, spmd_attach=None, n_gpus=1, allmax=lambda v: v, spmd_ctx=None,
                defer_finish=False):
    """C4: all-pairs n-body, 128 Ki bodies float4; pos is a whole-array read, vel/posOut are sliced."""
    n = 131072
    nb_src = kernels.source("nbody")
    if os.environ.get("CK_NBW_TPW"):        # experiment hook: j tiles per warp and unit (profiles/r2_nbody_unit_size.md)
        nb_src = f"// ck-nvrtc: -DNBW_TILES_PER_WARP={int(os.environ['CK_NBW_TPW'])}\n" + nb_src
    if os.environ.get("CK_NBW_LB"):         # experiment hook: "<max threads> <min groups per SM>" for nBody3Dw / nBody3Dwp
        nb_src = "".join(f"// ck-launch-bounds: {k} {os.environ['CK_NBW_LB']}\n" for k in ("nBody3Dw", "nBody3Dwp")) + nb_src
    cr = ck.ClNumberCruncher(devs, nb_src)
    assert cr.errorCode() == 0, cr.errorMessage()
    cr.balancerMode = 2
    if spmd_attach:
        spmd_attach(cr)
    pos = np.zeros((n, 4), np.float32)
    for c in range(3):
        pos[:, c] = po.uniform(0xC0FFEE + c, n, -15, 15)
    pos[:, 3] = 1.0
    # nBody3Dw: four work-items per body, work-groups of 128 pull (128 bodies x 8 tiles) units from a device-side queue
    kernel, per_body, local = "nBody3Dw", 4, 128
    groups, chunks = n // 128 + 8, (n // 256 + 7) // 8
    kfloats = (groups * chunks * 384 + groups + 256 + 1023) // 1024
    pa, pb, vel, params, scratch = (ck.ClFloatArray(4 * n), ck.ClFloatArray(4 * n), ck.ClFloatArray(4 * n), ck.ClFloatArray(8),
                                    ck.ClFloatArray(kfloats * 1024))
    pa.view[:] = pos.reshape(-1)
    params.view[:5] = [1e-3, 1e-4, n, per_body, kfloats]
    params.readOnly = True
    scratch.read, scratch.write = False, False          # device-resident partial sums and queue counters
    vel.partialRead, vel.read = True, False
    for a in (pa, pb, vel):
        a.numberOfElementsPerWorkItem = 4 // per_body
    src, dst = pa, pb
    ts_wall, ts_dev = [], []
    for i in range(warmup + steps):
        src.read, src.partialRead, src.write = True, False, False
        dst.read, dst.partialRead, dst.write = False, False, True
        t0 = time.perf_counter()
        assert src.nextParam(vel, dst, params, scratch).compute(cr, 1, kernel, n * per_body, local) == 0
        if i >= warmup:
            ts_wall.append(time.perf_counter() - t0)
            ts_dev.append(max(cr.benchmarks(1)) * 1e-3)
        src, dst = dst, src
    dt_dev, dt_wall = allmax(float(np.mean(ts_dev))), allmax(float(np.mean(ts_wall)))   # slowest rank decides
    # one more step, checked against one oracle step from the state this process held before it: 1,024 bodies of its own
    # share(s) against all n (with one process per GPU the other ranks' slices of the host copy are stale; the kernel and the
    # oracle both see that same input)
    before_p, before_v = src.view.copy(), vel.view.copy()
    src.read, src.partialRead, src.write = True, False, False
    dst.read, dst.partialRead, dst.write = False, False, True
    assert src.nextParam(vel, dst, params, scratch).compute(cr, 1, kernel, n * per_body, local) == 0
    parity = nbody_parity(po, before_p, before_v, dst.view, vel.view, n, own_bodies(cr, 1, per_body, spmd_ctx))
    src, dst = dst, src
    # the kernel alone: every array left on the device (flags cleared as Kamera.cs:271-272 does), shares as the balancer left
    # them.  With more than one device the other devices' position slices go stale, which changes no timing: the work per
    # body does not depend on the data.
    for a in (pa, pb, vel, params):
        a.read, a.partialRead, a.write = False, False, False
    ts_k = []
    for i in range(2 + steps):
        assert src.nextParam(vel, dst, params, scratch).compute(cr, 1, kernel, n * per_body, local) == 0
        if i >= 2:
            ts_k.append(max(cr.benchmarks(1)) * 1e-3)
        src, dst = dst, src
    dt_k = allmax(float(np.mean(ts_k)))
    inter = float(n) * n
    peak = 2.0 * 128 * sms * pk["sm_max_mhz"] * 1e6 / 1e12
    ach = 20.0 * inter / n_gpus / dt_k / 1e12   # per-GPU rate: the devices work on their shares concurrently
    res = {"bodies": n, "kernel": kernel, "ms_per_step_kernel": dt_k * 1e3, "ms_per_step_device": dt_dev * 1e3,
           "ms_per_step_e2e": dt_wall * 1e3, "GInter_per_s_kernel": inter / dt_k / 1e9,
           "GInter_per_s": inter / dt_dev / 1e9, "GInter_per_s_e2e": inter / dt_wall / 1e9, "scratch_MB": kfloats * 4096 / 1e6,
           "parity": parity,
           "roofline": {"bound": "fp32", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                        "traffic": ncu_traffic("nbody", "r2_ncu_summary.json") if n_gpus == 1 else None,
                        "note": "20 FLOP per interaction over the kernel's own launch time (arrays resident); GInter_per_s "
                                "adds the 2 MiB position upload and the slice transfers, GInter_per_s_e2e the whole host call"}}
    def finish():
        """The device-resident leg (when there is more than one device) and the clean-up.  Deferred on request so that the
        caller can have the rest of its result line ready before it starts (see the watchdog in run_ours)."""
        if n_gpus > 1 and (spmd_attach is None or spmd_ctx is not None):
            # Correct device-resident stepping with the exchange fused into the kernel (nBody3Dwp: every device stores its new
            # positions and velocities into all devices' buffers over NVLink — directly when one process drives all devices,
            # through CUDA IPC mappings when each GPU has its own process).  Positions go up once; after that a step is the
            # kernel and nothing else.  In one-process-per-GPU mode the exchange of device times that opens every compute()
            # is the barrier between steps.
            if spmd_ctx is not None:
                from cekirdekler_b200 import spmd as _spmd

                tabs = [_spmd.PeerTable(cr, spmd_ctx[0], spmd_ctx[1], spmd_ctx[2], x, vel) for x in (pa, pb)]
                t_a, t_b = tabs[0].array, tabs[1].array
                usable = all(t.ok for t in tabs)      # the same on every rank
            else:
                tabs = [cr.peerTable(pa, vel), cr.peerTable(pb, vel)]
                t_a, t_b = tabs
                usable = True
            pa.view[:] = pos.reshape(-1)
            params.readOnly = True
            src, dst, ts_p, ts_pw = pa, pb, [], []
            for i in range(3 + steps if usable else 0):
                src.read, src.partialRead, src.write = (i == 0), False, False
                t0 = time.perf_counter()
                assert src.nextParam(vel, dst, params, scratch, t_b if dst is pb else t_a).compute(cr, 2, "nBody3Dwp", n * per_body, local) == 0
                if i >= 3:
                    ts_pw.append(time.perf_counter() - t0)
                    ts_p.append(max(cr.benchmarks(2)) * 1e-3)
                src, dst = dst, src
            parity_p = None
            if usable and spmd_ctx is None:
                # two more steps that also download (untimed): the state before the last one and the state after it
                states = []
                for _ in range(2):
                    dst.write, vel.write = True, True
                    assert src.nextParam(vel, dst, params, scratch, t_b if dst is pb else t_a).compute(cr, 2, "nBody3Dwp", n * per_body, local) == 0
                    states.append((dst.view.copy(), vel.view.copy()))
                    dst.write, vel.write = False, False
                    src, dst = dst, src
                parity_p = nbody_parity(po, states[0][0], states[0][1], states[1][0], states[1][1], n, own_bodies(cr, 2, per_body, None))
            if usable:
                dt_p, dt_pw = allmax(float(np.mean(ts_p))), allmax(float(np.mean(ts_pw)))
                res["resident_peer_stores"] = {"kernel": "nBody3Dwp", "ms_per_step_device": dt_p * 1e3, "ms_per_step_e2e": dt_pw * 1e3,
                                               "GInter_per_s": inter / dt_p / 1e9, "GInter_per_s_e2e": inter / dt_pw / 1e9,
                                               "frac_of_fp32_per_gpu": 20.0 * inter / n_gpus / dt_p / 1e12 / peak,
                                               "exchange": "CUDA IPC peer stores" if spmd_ctx is not None else "peer stores",
                                               "parity": parity_p}
            else:
                res["resident_peer_stores"] = {"unavailable": "; ".join(t.error for t in tabs if t.error) or "a peer rank could not map the buffers"}
            for t in tabs:
                t.close() if spmd_ctx is not None else t.dispose()
        for a in (pa, pb, vel, params, scratch):
            a.dispose()
        cr.dispose()
        return res

    if defer_finish:
        return res, finish
    finish()
    return res


# ------------------------------------------------------------------------------------------------ GPU arm
def run_ours(args):
    rank, world, local = dist_env()
    import cekirdekler_b200 as ck
    from cekirdekler_b200 import kernels, spmd
    from cekirdekler_b200._native import host, shim
    from oracle import pyoracle as po  # checker + cpu_baseline leg only

    po.use_all_cores()   # torchrun exports OMP_NUM_THREADS=1; the CPU baseline leg states the cores it really used
    L = shim()
    if L.ckDeviceCount() <= 0:
        raise SystemExit("bench.py: no CUDA device visible; there is no CPU fallback (use --impl reference for the CPU arm)")
    pk = peaks()
    spmd_mode = world > 1
    numa = None
    if spmd_mode:
        spmd.init_process_group_from_env()
        if os.environ.get("CK_NUMA", "1") != "0":
            numa = spmd.prefer_host_memory_near_gpu(local)   # pinned arrays on the GPU's own socket
        import torch
        import torch.distributed as dist

        torch_dev = torch.device("cuda", local)
        devs = ck.ClDevices([local])
        n_gpus = world
    else:
        dist = None
        n_gpus = max(1, args.gpus)
        avail = L.ckDeviceCount()
        if n_gpus > avail:
            raise SystemExit(f"bench.py: --gpus {n_gpus} but only {avail} visible")
        devs = ck.ClDevices(list(range(n_gpus)))
    sms = host().ckh_gpu_compute_units(devs.ordinals[0])

    def barrier():
        if spmd_mode:
            dist.barrier()

    def allmax(v: float) -> float:
        if not spmd_mode:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=torch_dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allsum(v: float) -> float:
        if not spmd_mode:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=torch_dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    cr = ck.ClNumberCruncher(devs, kernels.source("mandelbrot"))
    assert cr.errorCode() == 0, cr.errorMessage()
    cr.balancerMode = {"reference": 0, "adaptive": 1, "profile": 2}[args.balancer]
    if args.no_smooth:
        cr.smoothLoadBalancer = False            # ClNumberCruncher.smoothLoadBalancer (ClNumberCruncher.cs:151-159)
    if spmd_mode:
        spmd.attach(cr, rank, world, torch_dev)
    npix = W * H
    view = ck.ClArray(8, np.float32)
    view.view[:] = po.mandelbrot_view(W, H, MAX_ITER)
    view.readOnly = True
    out = ck.ClArray(npix, np.int32)   # pinned host array, 1 GiB; every rank writes only its own slice
    out.writeOnly = True
    out.numberOfElementsPerWorkItem = E
    items = npix // E

    def my_slices(cid):
        r, f = cr.globalRanges(cid), cr.globalReferences(cid)
        idx = [rank] if spmd_mode else range(len(r))
        return [(f[i] * E, r[i] * E) for i in idx]

    # ---- leg 1: device-resident (value + roofline): no download, kernel span timed with CUDA events in the shim ----
    CID_RES = 1
    rw_res = [" read  ro ", ""]           # view: 32-byte upload; out: stays in HBM
    g = view.nextParam(out)
    shares, imbalance = [], []

    def step_resident():
        assert g.compute(cr, CID_RES, MK, items, LOCAL, 0, False, True, 4, readWritesReady=rw_res, elementsPerItemReady=[1, E]) == 0
        t_local = cr.benchmarks(CID_RES)
        return t_local

    for _ in range(args.warmup):
        t = step_resident()
        tt = cr_times(cr, CID_RES, t, spmd_mode, allgather=(lambda v: gather(dist, torch, torch_dev, v, world)) if spmd_mode else None)
        shares.append(cr.globalRanges(CID_RES))
        imbalance.append(max(tt) / (sum(tt) / len(tt)))
    # Multi-GPU: the iterative balancer is part of the path, and its convergence is reported below (iterations_to_converge).
    # The throughput figure is the steady state, so extra UNTIMED calls run until the split has settled (max/mean device
    # time <= 1.03 twice in a row, at most 16 calls); their number is disclosed in config.balancer_settle_steps.
    settle = 0
    while n_gpus > 1 and settle < args.max_settle and not (len(imbalance) >= 2 and max(imbalance[-2:]) <= 1.03):
        t = step_resident()
        tt = cr_times(cr, CID_RES, t, spmd_mode, allgather=(lambda v: gather(dist, torch, torch_dev, v, world)) if spmd_mode else None)
        shares.append(cr.globalRanges(CID_RES))
        imbalance.append(max(tt) / (sum(tt) / len(tt)))
        settle += 1
    barrier()
    cr.sync()
    launches0 = L.ckLaunchCount()
    dev_ms = 0.0
    with ClockSampler(devs.ordinals[0]) as clocks:
        t0 = time.perf_counter()
        for _ in range(args.steps):
            t = step_resident()
            tt = cr_times(cr, CID_RES, t, spmd_mode, allgather=(lambda v: gather(dist, torch, torch_dev, v, world)) if spmd_mode else None)
            dev_ms += max(tt)                  # devices run concurrently: the step takes as long as its slowest device
            shares.append(cr.globalRanges(CID_RES))
            imbalance.append(max(tt) / (sum(tt) / len(tt)))
        cr.sync()
        barrier()
        wall = time.perf_counter() - t0
    launches = L.ckLaunchCount() - launches0
    dev_ms = allmax(dev_ms)
    wall = allmax(wall)
    launches = int(allsum(float(launches)))
    ms_per_step = dev_ms / args.steps
    value = npix / (ms_per_step * 1e-3) / 1e6

    # ---- leg 2: end to end through the public API with host buffers (view upload + slice download every step) ----
    # Its own compute id: with host buffers a device's time is max(its kernels, the download of its slice), so the balanced
    # split is a different one (devices holding the cheap rows become host-link bound).  Settled like the resident leg.
    CID_E2E = 2
    e2e_t, e2e_imb, e2e_settle = [], [], 0

    def step_e2e():
        barrier()
        t0 = time.perf_counter()
        assert g.compute(cr, CID_E2E, MK, items, LOCAL, 0, True, args.e2e_mode == "event", args.e2e_blobs) == 0
        dt = allmax(time.perf_counter() - t0)
        tt = cr_times(cr, CID_E2E, cr.benchmarks(CID_E2E), spmd_mode,
                      allgather=(lambda v: gather(dist, torch, torch_dev, v, world)) if spmd_mode else None)
        e2e_imb.append(max(tt) / (sum(tt) / len(tt)))
        return dt

    for _ in range(args.warmup):
        step_e2e()
    while n_gpus > 1 and e2e_settle < args.max_settle and not (len(e2e_imb) >= 2 and max(e2e_imb[-2:]) <= 1.03):
        step_e2e()
        e2e_settle += 1
    for _ in range(args.steps):
        e2e_t.append(step_e2e())
    e2e_ms = float(np.mean(e2e_t)) * 1e3
    e2e_value = npix / (e2e_ms * 1e-3) / 1e6

    # ---- work accounting and in-run parity on the oracle's strided sample ----
    local_iters = 0
    for (p0, cnt) in my_slices(CID_E2E):
        local_iters += int(out.view[p0:p0 + cnt].sum(dtype=np.int64))
    total_iters = allsum(float(local_iters))
    line_cpu, parity = None, None
    # every rank checks its OWN slice of the image against the oracle on a stride-257 sample (the ranks' shares differ, and
    # rank 0's larger sample below only sees rank 0's); mismatches and pixel counts are summed over the ranks
    own_checked, own_bad = 0, 0
    if spmd_mode:
        po.lib().oracle_set_threads(max(1, po.threads() // max(1, world)))   # the ranks share the host's cores here
        for (p0, cnt) in my_slices(CID_E2E):
            m = (cnt - 17) // 257
            if m <= 0:
                continue
            ref = np.zeros(m, np.int32)
            po.mandelbrot_sample(view.view, p0 + 17, 257, m, ref)
            got = out.view[p0 + 17:p0 + 17 + 257 * m:257]
            own_checked += m
            own_bad += int((got != ref).sum())
        po.use_all_cores()
    all_checked, all_bad = int(allsum(float(own_checked))), int(allsum(float(own_bad)))
    if rank == 0:
        s = cpu_sample(po, view.view, 12.0)
        idx = s["offset"] + s["stride"] * np.arange(s["pixels"], dtype=np.int64)
        mine = np.zeros(npix, bool)
        for (p0, cnt) in my_slices(CID_E2E):
            mine[p0:p0 + cnt] = True
        sel = mine[idx]
        parity = {"sample_pixels_checked": int(sel.sum()), "mismatches": int((out.view[idx[sel]] != s["values"][sel]).sum())}
        if spmd_mode:
            parity["every_rank_own_slice"] = {"sample_pixels_checked": all_checked, "mismatches": all_bad, "stride": 257}
        line_cpu = {"value": s["pixels"] / s["seconds"] / 1e6, "unit": UNIT, "cores": po.threads(), "kind": "port",
                    "sample": f"{s['pixels']} pixels (odd stride {s['stride']} across the whole {W}x{H} image), {s['iters']} iterations, {s['seconds']:.1f} s"}

    flops = 8.0 * total_iters
    peak_tf = 2.0 * 128 * sms * pk["sm_max_mhz"] * 1e6 / 1e12
    kernel_s = dev_ms / args.steps * 1e-3      # resident leg: the span is the kernel (+ a 32-byte upload)
    ach = flops / n_gpus / kernel_s / 1e12     # per-GPU rate (each GPU runs its share concurrently)
    clk = clocks.summary()
    roof = {"bound": "fp32", "achieved": ach, "peak": peak_tf, "unit": "TFLOP/s", "frac": ach / peak_tf,
            "traffic": (ncu_traffic("mandelbrot1", "r2_ncu_summary.json") if MK == "mandelbrot1" else ncu_traffic("mandelbrot")) if n_gpus == 1 else None,
            "traffic_note": "DRAM read+write bytes of one launch (ncu --set full, profiles/r2_ncu_summary.json / r1_ncu_summary.json); the algorithmic "
                            "output is 4 B x 2^28 pixels = 1.074e9 B",
            "kernel": MK, "per_gpu": True,
            "algorithmic": f"8 FLOP x {int(total_iters)} executed iterations (sum of the output counts, bit-exact with the oracle)",
            "peak_source": f"2 x 128 lanes x {sms} SMs x {pk['sm_max_mhz']:.0f} MHz ({pk['source']}); no tensor cores on this path",
            "frac_at_observed_clock": (ach / (2.0 * 128 * sms * clk["sm_mhz"] * 1e6 / 1e12)) if clk.get("sm_mhz") else None}

    workloads = {}
    link_all = None
    if args.all:
        # every GPU of the run copies at the same time: the ceiling the end-to-end leg can be held against
        if spmd_mode:
            mine = bench_hostlink(L, (devs.ordinals[0],), 256, 4, barrier=barrier)
            link_all = {k: allsum(v) for k, v in mine.items() if k.endswith("_GBps")}
            link_all["per_rank_d2h_GBps"] = gather(dist, torch, torch_dev, mine["d2h_GBps"], world)
            link_all["how"] = f"{world} ranks, one GPU each, 256 MiB pinned copies started together (barrier), sum over ranks"
        elif n_gpus > 1:
            link_all = bench_hostlink(L, tuple(devs.ordinals), 256, 4)
            link_all["how"] = f"one process, {n_gpus} GPUs, 256 MiB pinned copies enqueued on all devices, then drained"
    if args.all and rank == 0 and not spmd_mode:
        link = bench_hostlink(L, (devs.ordinals[0],))
        workloads["host_link"] = link
        sx = bench_saxpy(ck, kernels, devs[0], L, pk, max(2, args.steps // 2), 1)
        for mode in ("streamed_event", "streamed_driver"):
            # streamed roofline: the upload direction carries 8 of the 12 bytes per element and is the binding one
            sx[mode]["roofline"] = {"bound": "host_link", "achieved": sx[mode]["h2d_GBps"], "peak": link["bidirectional_each_GBps"],
                                    "unit": "GB/s", "frac": sx[mode]["h2d_GBps"] / link["bidirectional_each_GBps"],
                                    "peak_source": "pinned H2D with D2H running concurrently, measured in this run"}
        workloads["saxpy_c2"] = sx
        try:
            workloads["mandelbrot_c1"] = bench_c1(ck, kernels, po, devs[0], pk, sms)
        except Exception as e:
            workloads["mandelbrot_c1"] = {"error": repr(e)}
        workloads["pipeline_c5"] = bench_pipeline(ck, kernels, po, n_gpus)
        try:
            workloads["pipeline_c5"]["kernels_resident"] = bench_image_kernels(ck, kernels, po, devs[0], pk)
        except Exception as e:   # an extra leg must not cost the line
            workloads["pipeline_c5"]["kernels_resident"] = {"error": repr(e)}
    if args.all and spmd_mode and world >= 3:
        # C5 with its three stages on GPUs 0, 1 and 2: rank 0 drives them from its own process while the other ranks wait on
        # the host (a shared-memory rendezvous, no kernel of theirs spins on the GPUs that are being measured)
        hb = spmd.make_exchange(rank, world, torch_dev)
        if rank == 0:
            try:
                workloads["pipeline_c5_3gpu"] = bench_pipeline(ck, kernels, po, 3)
            except Exception as e:
                workloads["pipeline_c5_3gpu"] = {"error": repr(e)}
        hb([0.0] * world)
    if args.all:
        attach = (lambda c: spmd.attach(c, rank, world, torch_dev)) if spmd_mode else None
        nb, nbody_finish = bench_nbody(ck, kernels, po, devs, pk, sms, max(3, args.steps), 2, attach, n_gpus, allmax,
                                       (rank, world, torch_dev) if spmd_mode else None, defer_finish=True)
        workloads["nbody_c4"] = nb
    else:
        nbody_finish = None

    line = None
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n_gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD,
                       "kernel": f"{MK}, {E} pixel(s) per work-item, local {LOCAL}, via ClNumberCruncher.compute",
                       "parallelism": ("spmd one process per GPU" if spmd_mode else "one process") + f" x{n_gpus}",
                       "l2": "output 1 GiB per step >> 126 MB L2, no reuse between steps",
                       "balancer": args.balancer, "smooth_load_balancer": not args.no_smooth, "balancer_settle_steps": settle, "host_memory_policy_rank0": numa},
            "wall_ms_per_step": wall / args.steps * 1e3,
            "clocks": clk,
            "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": e2e_ms, "h2d_bytes_per_step": 32 * n_gpus,
                    "d2h_bytes_per_step": 4 * npix, "mode": f"{args.e2e_mode}-controlled pipeline, {args.e2e_blobs} blobs, pinned host arrays",
                    "balancer_settle_steps": e2e_settle,
                    "balancer": {"iterations_to_converge": first_stable(e2e_imb) if n_gpus > 1 else 0,
                                 "imbalance_max_over_mean": [round(v, 4) for v in e2e_imb],
                                 "final_shares": [round(x / items, 4) for x in cr.globalRanges(CID_E2E)]}},
            "gpu_launches": launches,
            "roofline": roof,
            "cpu_baseline": line_cpu,
            "parity": parity,
            "balancer": {"iterations_to_converge": first_stable(imbalance) if n_gpus > 1 else 0,
                         "imbalance_max_over_mean": [round(v, 4) for v in imbalance],
                         "final_shares": [round(x / items, 4) for x in shares[-1]] if shares else None},
            "workloads": workloads,
        }
        if link_all is not None:
            workloads["host_link_all_gpus"] = link_all
            d2h = link_all.get("d2h_GBps")
            if d2h:
                line["e2e"]["d2h_GBps_achieved"] = 4.0 * npix / (e2e_ms * 1e-3) / 1e9
                line["e2e"]["d2h_GBps_ceiling_all_gpus"] = d2h
                line["e2e"]["frac_of_host_link_ceiling"] = line["e2e"]["d2h_GBps_achieved"] / d2h
    # The last leg — n-body stepping with the exchange fused into the kernel, which at N > 1 maps other ranks' buffers and
    # relies on every rank taking part — runs with the rest of the line already complete.  If it has not finished within
    # 120 s (a rank that hangs would otherwise take the whole measurement with it), rank 0 prints the line without it and
    # every rank leaves.
    finished = threading.Event()

    def watchdog():
        if not finished.wait(120.0):
            if line is not None:
                workloads.get("nbody_c4", {})["resident_peer_stores"] = {"unavailable": "no result within 120 s"}
                emit(line)
            os._exit(0)

    if nbody_finish is not None:
        threading.Thread(target=watchdog, daemon=True).start()
        nbody_finish()
    finished.set()
    if line is not None:
        emit(line)
    if spmd_mode:
        dist.barrier()
        dist.destroy_process_group()


def gather(dist, torch, dev, v: float, world: int):
    mine = torch.tensor([v], dtype=torch.float64, device=dev)
    outt = torch.empty(world, dtype=torch.float64, device=dev)
    dist.all_gather_into_tensor(outt, mine)
    return outt.cpu().tolist()


def cr_times(cr, cid, local_times, spmd_mode, allgather=None):
    """Per-device times of the last call over ALL devices of the split."""
    if not spmd_mode:
        return [max(t, 1e-9) for t in local_times]
    return [max(t, 1e-9) for t in allgather(local_times[0])]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--balancer", default="profile", choices=["profile", "adaptive", "reference"])
    ap.add_argument("--mandel-kernel", default="mandelbrot1", choices=["mandelbrot1", "mandelbrot"],
                    help="mandelbrot1: one pixel per work-item (round 2); mandelbrot: four pixels of a row per work-item (round 1)")
    ap.add_argument("--no-smooth", action="store_true", help="smoothLoadBalancer = false (the reference's default is true)")
    ap.add_argument("--max-settle", type=int, default=16, help="untimed balancer calls allowed before the timed steps (0: time from the start)")
    ap.add_argument("--no-all", dest="all", action="store_false", help="skip the SAXPY and n-body workloads")
    ap.add_argument("--e2e-blobs", type=int, default=16, help="pipeline blobs of the end-to-end leg (multiple of 4)")
    ap.add_argument("--e2e-mode", default="event", choices=["event", "driver"], help="pipelining mode of the end-to-end leg")
    ap.add_argument("--only", default=None, choices=["nbody", "saxpy", "pipeline", "hostlink", "c1", "image"],
                    help="profiling aid: run just this secondary workload on one GPU and print its numbers")
    args = ap.parse_args()
    global E, MK
    MK, E = args.mandel_kernel, (4 if args.mandel_kernel == "mandelbrot" else 1)
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    # stdout carries exactly ONE JSON line: libraries that print banners there (NCCL's version line) go to stderr
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.only:
        import cekirdekler_b200 as ck
        from cekirdekler_b200 import kernels
        from cekirdekler_b200._native import host, shim
        from oracle import pyoracle as po

        dev = ck.ClPlatforms.all().gpus()[0]
        pk = peaks()
        if args.only == "nbody":
            out = bench_nbody(ck, kernels, po, dev, pk, host().ckh_gpu_compute_units(0), args.steps, args.warmup)
        elif args.only == "pipeline":
            out = bench_pipeline(ck, kernels, po, max(1, args.gpus))
            out["kernels_resident"] = bench_image_kernels(ck, kernels, po, dev, pk)
        elif args.only == "hostlink":
            out = bench_hostlink(shim(), tuple(range(max(1, args.gpus))))
        elif args.only == "c1":
            out = bench_c1(ck, kernels, po, dev, pk, host().ckh_gpu_compute_units(0))
        elif args.only == "image":
            out = bench_image_kernels(ck, kernels, po, dev, pk)
        else:
            out = bench_saxpy(ck, kernels, dev, shim(), pk, args.steps, args.warmup)
        emit({"workload": args.only, **out})
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
