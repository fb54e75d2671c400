"""Does a GPU that sat idle for a few milliseconds (as a rank does at the barrier in front of a timed region while
it waits for the slowest rank) run its next 20 pipelined steps slower?  One GPU: the bench's pipelined dense loop
(64 frames, three contexts in flight), 20 steps, CUDA-event timed, after an idle gap of 0 ... 200 ms, 12 repeats each.
usage: python tools/idle_gap_probe.py"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfpose_b200 as tp
from tfpose_b200 import synth

B, h, w = 64, 46, 54
H, W = 4 * h, 4 * w
xs = []
for s in range(4):
    base = synth.synth_batch(8, h, w, config_id=2, first_frame=8 * s)
    xs.append(torch.from_numpy(np.concatenate([base] * 8)).cuda())
n_pipe = 3
decs = [tp.PoseDecoder(max_batch=B, max_upsample=(H, W)) for _ in range(n_pipe)]
denses = [tuple(torch.empty((B, H, W, c), device="cuda") for c in (19, 38, 19)) for _ in range(n_pipe)]
streams = [torch.cuda.current_stream()] + [torch.cuda.Stream() for _ in range(n_pipe - 1)]
for d in decs:
    d.set_option("k1_ctas", 128)
    d.set_option("k1_turnstile", 1)


def run(n):
    inflight = []
    for i in range(n):
        j = i % n_pipe
        if len(inflight) == n_pipe:
            inflight.pop(0).fetch(reuse=True)
        decs[j].decode_async(xs[i % 4], (H, W), dense=denses[j], stream=streams[j])
        inflight.append(decs[j])
    for d in inflight:
        d.fetch(reuse=True)
    for st in streams[1:]:
        streams[0].wait_stream(st)


e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for gap_ms in (0, 1, 3, 10, 30, 100, 200, 0):
    t = []
    for rep in range(12):
        run(5)                                  # warm-up
        torch.cuda.synchronize()
        if gap_ms:
            time.sleep(gap_ms * 1e-3)
        e0.record()
        run(20)
        e1.record()
        torch.cuda.synchronize()
        t.append(e0.elapsed_time(e1) / 20)
    t = np.array(t)
    print("idle gap %4d ms: ms per step median %.4f  min %.4f  max %.4f   %s" % (gap_ms, np.median(t), t.min(), t.max(), np.round(t, 4).tolist()), flush=True)
for d in decs:
    d.close()
