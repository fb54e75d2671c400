"""Where a step's GPU time goes when steps run back to back on one stream (no fetch in between): event-timed
stages of the last step next to the measured time per step.  usage: python tools/chain_times.py"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfpose_b200 as tp
from tfpose_b200 import synth

h, w, B, H, W = 46, 54, 64, 184, 216
xs = [torch.from_numpy(synth.synth_batch(B, h, w, config_id=2, first_frame=64 * i)).cuda() for i in range(4)]
st = torch.cuda.Stream()
for ctas in (0, 128):
    for prof in (True, False):
        for graph in (0, 1):
            if prof and graph:
                continue
            dec = tp.PoseDecoder(max_batch=B, max_upsample=(H, W))
            dec.set_profiling(prof)
            dec.set_option("k1_ctas", ctas)
            dec.set_option("graph", graph)
            for mode in ("minimal", "dense"):
                dense = tuple(torch.empty((B, H, W, c), device="cuda") for c in (19, 38, 19)) if mode == "dense" else None
                for i in range(8):
                    dec.decode_async(xs[i % 4], (H, W), dense=dense, stream=st)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                R = 40
                with torch.cuda.stream(st):
                    e0.record()
                    for i in range(R):
                        dec.decode_async(xs[i % 4], (H, W), dense=dense, stream=st)
                    e1.record()
                torch.cuda.synchronize()
                per = e0.elapsed_time(e1) / R * 1e3
                stages = np.array(dec.stage_ms()) * 1e3 if prof else None
                dec.fetch()
                t0 = time.perf_counter()
                for i in range(R):
                    dec.decode(xs[i % 4], (H, W), dense=dense, stream=st)
                sync_us = (time.perf_counter() - t0) / R * 1e6
                print("k1_ctas %3d profiling %-5s graph %d %-7s %.1f us/step back to back, %.1f us per decode+fetch%s"
                      % (ctas, prof, graph, mode, per, sync_us,
                         ("  stages %s sum %.1f" % (np.round(stages, 1), stages.sum())) if prof else ""))
            dec.close()
