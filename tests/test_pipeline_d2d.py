"""Device-to-device pipeline (ClPipeline.cs:41-2356): blur -> Sobel -> histogram over a stream of frames, one stage per
device (all on device 0 when the box has one GPU), peer-to-peer hand-off vs the reference's host bounce.  Both must
deliver, for every frame, exactly the oracle's histogram, with the reference's latency (results valid once pushData
returns true: counter > 2S with data and popResultsHere both given, ClPipeline.cs:114-124)."""
import numpy as np
import pytest

import cekirdekler_b200 as ck
from cekirdekler_b200 import kernels
from cekirdekler_b200.pipeline import ClPipelineStage
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("host_forwarding", [False, True], ids=["p2p", "host"])
def test_three_stage_pipeline_histograms_are_exact(gpus, host_forwarding):
    W, H = 512, 256
    n = W * H
    src = kernels.source("image") + """
    // stage wrappers: parameter order is (inputs..., outputs...) as ClPipelineStage builds it (ClPipeline.cs:281-370)
    __kernel void blurStage(const float* in, const int* dims, float* out, int* dimsOut) {
        const size_t p = get_global_id(0);
        if (p < 2) dimsOut[p] = dims[p];
        const int W = dims[0], H = dims[1];
        if (W <= 0 || H <= 0) return;     // pipeline warm-up: no frame has arrived at this stage yet
        const int y = (int)(p / (unsigned)W), x = (int)(p - (size_t)y * W);
        const float c = ck_blur_h(in, W, x, y);
        const float s1 = __fadd_rn(ck_blur_h(in, W, x, ck_clampi(y - 1, 0, H - 1)), ck_blur_h(in, W, x, ck_clampi(y + 1, 0, H - 1)));
        const float s2 = __fadd_rn(ck_blur_h(in, W, x, ck_clampi(y - 2, 0, H - 1)), ck_blur_h(in, W, x, ck_clampi(y + 2, 0, H - 1)));
        out[p] = __fmul_rn(0.00390625f, __fmaf_rn(6.0f, c, __fmaf_rn(4.0f, s1, s2)));
    }
    """
    devs = ck.ClPlatforms.all().gpus()
    d = lambda i: devs[i % gpus]

    blur = ClPipelineStage()
    blur.addDevices(d(0))
    blur.addKernels(src, "blurStage", [n], [256])
    blur.addInputBuffers((np.float32, n), (np.int32, 2))
    blur.addOutputBuffers((np.float32, n), (np.int32, 2))
    sob = ClPipelineStage()
    sob.addDevices(d(1))
    # sobel(in, out, dims) expects dims last: inputs are (frame, dims), so wrap the order with a hidden-free stage kernel
    sob.addKernels(src + """
    __kernel void sobelStage(const float* in, const int* dims, float* out) {
        const int W = dims[0], H = dims[1];
        if (W <= 0 || H <= 0) return;
        const size_t p = get_global_id(0);
        const int y = (int)(p / (unsigned)W), x = (int)(p - (size_t)y * W);
        const int xm = ck_clampi(x - 1, 0, W - 1), xp = ck_clampi(x + 1, 0, W - 1);
        const int ym = ck_clampi(y - 1, 0, H - 1), yp = ck_clampi(y + 1, 0, H - 1);
        const float* rm = in + (size_t)ym * W; const float* r0 = in + (size_t)y * W; const float* rp = in + (size_t)yp * W;
        const float right = __fmaf_rn(2.0f, r0[xp], __fadd_rn(rm[xp], rp[xp]));
        const float left = __fmaf_rn(2.0f, r0[xm], __fadd_rn(rm[xm], rp[xm]));
        const float bottom = __fmaf_rn(2.0f, rp[x], __fadd_rn(rp[xm], rp[xp]));
        const float top = __fmaf_rn(2.0f, rm[x], __fadd_rn(rm[xm], rm[xp]));
        const float gx = __fsub_rn(right, left), gy = __fsub_rn(bottom, top);
        out[p] = __fsqrt_rn(__fmaf_rn(gx, gx, __fmul_rn(gy, gy)));
    }""", "sobelStage", [n], [256])
    sob.addInputBuffers((np.float32, n), (np.int32, 2))
    sob.addOutputBuffers((np.float32, n))               # written one pixel per work-item
    sob.appendToStage(blur)
    hist = ClPipelineStage()
    hist.addDevices(d(2))
    hist.addKernels(src, "histogramZero histogram", [n // 16, n // 16], [256, 256])
    hist.addInputBuffers((np.float32, n, 16))
    hist.addOutputBuffers((np.uint32, 256))
    hist.appendToStage(sob)
    pipe = hist.makePipeline(hostForwarding=host_forwarding)
    assert pipe.errorCode() == 0, pipe.errorMessage()
    assert pipe.peerToPeer == (not host_forwarding)

    frames = [po.uniform(100 + f, n, 0, 1) for f in range(10)]
    dims = np.array([W, H], np.int32)
    expected = []
    for fr in frames:
        b, s, h = np.zeros(n, np.float32), np.zeros(n, np.float32), np.zeros(256, np.uint32)
        po.blur5x5(fr, b, dims)
        po.sobel(b, s, dims)
        po.histogram(s, h)
        expected.append(h)

    got = []
    result = np.zeros(256, np.uint32)
    S = 3
    pushes = 0
    feed = frames + [frames[-1]] * (2 * S + 1)          # keep pushing until the last real frame has drained
    for fr in feed:
        valid = pipe.pushData([fr, dims], [result])
        pushes += 1
        assert valid == (pushes > 2 * S)                 # ClPipeline.cs:122: counter > 2S with data and results given
        if valid:
            got.append(result.copy())
    assert len(got) >= len(frames)
    for f in range(len(frames)):
        assert np.array_equal(got[f], expected[f]), f"frame {f}"
        assert int(got[f].sum()) == n
