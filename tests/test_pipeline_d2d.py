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


def build_pipeline(d, W, H, rows, host_forwarding, fused=False):
    """blur -> Sobel -> histogram as three stages on d(0), d(1), d(2) (strip kernels), or the fused kernel as a single stage."""
    n = W * H
    src = kernels.image_pipeline_source()
    g, l = kernels.strip_launch(W, H, rows)
    if fused:
        one = ClPipelineStage()
        one.addDevices(d(0))
        one.addKernels(src, "fusedStage", [g], [l])
        one.addInputBuffers((np.float32, n), (np.int32, 3))
        one.addHiddenBuffers((np.uint32, 2049))
        one.addOutputBuffers((np.uint32, 256))
        return one.makePipeline(hostForwarding=host_forwarding), 1
    blur = ClPipelineStage()
    blur.addDevices(d(0))
    blur.addKernels(src, "blurStage", [g], [l])
    blur.addInputBuffers((np.float32, n), (np.int32, 3))
    blur.addOutputBuffers((np.float32, n), (np.int32, 3))
    sob = ClPipelineStage()
    sob.addDevices(d(1))
    sob.addKernels(src, "sobelStage", [g], [l])
    sob.addInputBuffers((np.float32, n), (np.int32, 3))
    sob.addOutputBuffers((np.float32, n), (np.int32, 3))
    sob.appendToStage(blur)
    hist = ClPipelineStage()
    hist.addDevices(d(2))
    hist.addKernels(src, "histStage", [148 * 2 * 256], [256])
    hist.addInputBuffers((np.float32, n), (np.int32, 3))
    hist.addHiddenBuffers((np.uint32, 2049))            # the kernel's accumulator: zero when created, re-armed by every launch
    hist.addOutputBuffers((np.uint32, 256))
    hist.appendToStage(sob)
    return hist.makePipeline(hostForwarding=host_forwarding), 3


@pytest.mark.parametrize("fused", [False, True], ids=["three-stages", "fused-stage"])
@pytest.mark.parametrize("host_forwarding", [False, True], ids=["p2p", "host"])
def test_three_stage_pipeline_histograms_are_exact(gpus, host_forwarding, fused):
    W, H, rows = 512, 256, 16
    n = W * H
    devs = ck.ClPlatforms.all().gpus()
    d = lambda i: devs[i % gpus]
    pipe, S = build_pipeline(d, W, H, rows, host_forwarding, fused)
    assert pipe.errorCode() == 0, pipe.errorMessage()
    assert pipe.peerToPeer == (not host_forwarding)

    frames = [po.uniform(100 + f, n, 0, 1) for f in range(10)]
    dims = np.array([W, H, rows], np.int32)
    expected = []
    for fr in frames:
        b, s, h = np.zeros(n, np.float32), np.zeros(n, np.float32), np.zeros(256, np.uint32)
        po.blur5x5(fr, b, dims[:2].copy())
        po.sobel(b, s, dims[:2].copy())
        po.histogram(s, h)
        expected.append(h)

    got = []
    result = np.zeros(256, np.uint32)
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
