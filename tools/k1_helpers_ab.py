"""K1 dense with the two helper modes ("k1_helpers": 0 two double-slot helpers, 1 up to four single-slot ones), same
process, same box, rotating inputs, the BASELINE shapes; then the per-CTA role timeline of both modes on c2
(TFPD_K1_TIMES probe: when each CTA's blur role and copy warp finish).  usage: python tools/k1_helpers_ab.py"""
import os
import sys

import numpy as np
import torch

os.environ.setdefault("TFPD_K1_TIMES", "1")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfpose_b200 as tp
from tfpose_b200 import synth

for name, h, w, cfg in (("c2", 46, 54, 2), ("c3", 46, 82, 3), ("c5", 92, 164, 5)):
    B, H, W = 64, 4 * h, 4 * w
    xs = [torch.from_numpy(np.concatenate([synth.synth_batch(8, h, w, config_id=cfg, first_frame=8 * s)] * 8)).cuda() for s in range(4)]
    dense = tuple(torch.empty((B, H, W, c), device="cuda") for c in (19, 38, 19))
    dec = tp.PoseDecoder(max_batch=B, max_upsample=(H, W))
    dec.set_profiling(True)
    for ctas in (128, 148):
        dec.set_option("k1_ctas", ctas)
        for rep in range(2):
            for mode in (0, 1):
                dec.set_option("k1_helpers", mode)
                t = []
                for i in range(15):
                    dec.decode(xs[i % 4], (H, W), dense=dense)
                    if i >= 3:
                        t.append(dec.stage_ms()[0])
                line = "%s %d CTAs k1_helpers %d: %.1f us (min %.1f)" % (name, ctas, mode, np.mean(t) * 1e3, np.min(t) * 1e3)
                if name == "c2" and rep == 1:
                    tt = dec.k1_cta_times(ctas).astype(np.int64)
                    t0 = tt[:, 0].min()
                    blur, copy = (tt[:, 1] - t0) / 1e3, (tt[:, 2] - t0) / 1e3
                    line += " | blur role ends: median %.1f p90 %.1f max %.1f | dedicated copy warp ends: median %.1f max %.1f" % (
                        np.median(blur), np.percentile(blur, 90), blur.max(), np.median(copy), copy.max())
                print(line)
    dec.close()
    del dense, xs
    torch.cuda.empty_cache()
