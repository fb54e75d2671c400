"""Single-frame latency (BASELINE.json configs[0] shape, 46x54x57 -> 184x216): the reference's own use of the path
(TfPoseEstimator.inference decodes one frame per call).  Wall clock per call, device input and host input."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfpose_b200 as tp
from tfpose_b200 import synth

h, w = 46, 54
H, W = 4 * h, 4 * w
x = synth.synth_batch(1, h, w, config_id=2)
xd = torch.from_numpy(x).cuda()
xp = torch.from_numpy(x).pin_memory().numpy()
dec = tp.PoseDecoder(max_batch=1, max_upsample=(H, W))
dec.set_profiling(True)
dense = tuple(torch.empty((1, H, W, c), device="cuda") for c in (19, 38, 19))
post = tp.TfPosePostProcessor(target_size=(432, 368), keep_maps=True)


def timeit(fn, n=200, warm=20):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e6


print("decode, device input, minimal      %.1f us/frame" % timeit(lambda: dec.decode(xd, (H, W))))
print("   stages (ms):", ["%.4f" % v for v in dec.stage_ms()])
print("decode, device input, dense maps   %.1f us/frame" % timeit(lambda: dec.decode(xd, (H, W), dense=dense)))
print("   stages (ms):", ["%.4f" % v for v in dec.stage_ms()])
dec.set_profiling(False)
print("decode, minimal, profiling off     %.1f us/frame" % timeit(lambda: dec.decode(xd, (H, W))))
dec.set_option("graph", 1)
st = torch.cuda.Stream()
print("decode, minimal, CUDA graph replay %.1f us/frame" % timeit(lambda: dec.decode(xd, (H, W), stream=st)))
print("decode, dense, CUDA graph replay   %.1f us/frame" % timeit(lambda: dec.decode(xd, (H, W), dense=dense, stream=st)))
dec.set_option("graph", 0)
dec.set_profiling(True)
print("decode_host, pinned input, minimal %.1f us/frame" % timeit(lambda: dec.decode_host(xp, (H, W))))
print("decode + Human objects             %.1f us/frame" % timeit(lambda: dec.decode(xd, (H, W)).humans(0)))
print("inference_from_output (numpy in, heatMat/pafMat out like inference()) %.1f us/frame"
      % timeit(lambda: post.inference_from_output(x[0], upsample_size=4.0), n=50, warm=5))
post2 = tp.TfPosePostProcessor(target_size=(432, 368), keep_maps=True, reuse_map_buffers=True)
print("same, reuse_map_buffers=True (page-locked heatMat/pafMat)             %.1f us/frame"
      % timeit(lambda: post2.inference_from_output(x[0], upsample_size=4.0), n=50, warm=5))
