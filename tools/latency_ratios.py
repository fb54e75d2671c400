"""Single-frame decode latency per upsample ratio (46x54x57 network output), dense and minimal mode.
usage: python tools/latency_ratios.py"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfpose_b200 as tp
from tfpose_b200 import synth

h, w = 46, 54
x = torch.from_numpy(synth.synth_batch(1, h, w, config_id=2)).cuda()
for ratio in (1, 2, 3, 4, 8):
    H, W = h * ratio, w * ratio
    dec = tp.PoseDecoder(max_batch=1, max_upsample=(H, W), max_peaks=4096, max_cands=8192)
    dense = tuple(torch.empty((1, H, W, c), device="cuda") for c in (19, 38, 19))
    for mode, d in (("dense", dense), ("minimal", None)):
        dec.set_profiling(True)
        for _ in range(5):
            dec.decode(x, (H, W), dense=d)
        st = np.array(dec.stage_ms()) * 1e3
        dec.set_profiling(False)
        for _ in range(20):
            dec.decode(x, (H, W), dense=d)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(200):
            res = dec.decode(x, (H, W), dense=d)
        dt = (time.perf_counter() - t0) / 200 * 1e6
        print("ratio %d %-7s %.1f us per call | stages maps %.1f sort %.1f K2 %.1f K3 %.1f | %d humans, %d launches"
              % (ratio, mode, dt, st[0], st[1], st[2], st[3], int(res.info["n_humans"].sum()), dec.launch_count()))
    dec.close()
