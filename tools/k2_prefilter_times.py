"""Limb-scoring kernel (K2) with and without the early rejection ("k2_prefilter"), sparse (c2) and crowded (c4)
frames, 64 frames per call, minimal mode, K2 timed alone with CUDA events -- and against the peak-table capacity:
K2's shared memory is 10 B per peak slot, which sets how many of its CTAs share an SM.
usage: python tools/k2_prefilter_times.py"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfpose_b200 as tp
from tfpose_b200 import synth

B, h, w = 64, 46, 54
for name, cfg, people in (("c2 sparse", 2, None), ("c4 crowded", 4, 30)):
    base = synth.synth_batch(16, h, w, config_id=cfg, **({"people": people} if people else {}))
    x = torch.from_numpy(np.concatenate([base] * 4)).cuda()
    for dense_mode in (False, True):
      dense = tuple(torch.empty((B, 184, 216, c), device="cuda") for c in (19, 38, 19)) if dense_mode else None
      for caps in (dict(max_peaks=2048, max_cands=4096), dict(max_peaks=4096, max_cands=8192),
                   dict(max_peaks=4096, max_cands=8192, max_humans=128), dict(max_peaks=2048, max_cands=4096, max_humans=128)):
        dec = tp.PoseDecoder(max_batch=B, max_upsample=(184, 216), **caps)
        dec.set_profiling(True)
        for pre in (0, 1):
            dec.set_option("k2_prefilter", pre)
            t = []
            for i in range(12):
                res = dec.decode(x, (184, 216), dense=dense)
                if i >= 2:
                    t.append(dec.stage_ms())
            st = np.mean(t, 0) * 1e3
            print("%-10s %-7s %-58s k2_prefilter %d: K2 %.1f us (K1 %.1f, K3 %.1f) | candidates/frame %.0f"
                  % (name, "dense" if dense_mode else "minimal", caps, pre, st[2], st[0], st[3], res.info["n_cands"].mean()))
        dec.close()
