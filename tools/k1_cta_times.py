"""Per-CTA finish times of the fused map kernel on the bench workload (TFPD_K1_TIMES probe).
usage: TFPD_K1_TIMES=1 python tools/k1_cta_times.py"""
import os
import sys

import numpy as np
import torch

os.environ.setdefault("TFPD_K1_TIMES", "1")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfpose_b200 as tp
from tfpose_b200 import synth

B, h, w = 64, 46, 54
H, W = 4 * h, 4 * w
base = synth.synth_batch(16, h, w, config_id=2)
x = torch.from_numpy(np.concatenate([base] * 4)).cuda()
dense = (torch.empty((B, H, W, 19), device="cuda"), torch.empty((B, H, W, 38), device="cuda"),
         torch.empty((B, H, W, 19), device="cuda"))
dec = tp.PoseDecoder(max_batch=B, max_upsample=(H, W))
dec.set_profiling(True)
NC = int(os.environ.get("K1_CTAS", "128"))
dec.set_option("k1_ctas", NC)
for it in range(6):
    dec.decode(x, (H, W), dense=dense)
    t = dec.k1_cta_times(NC).astype(np.int64)
    t0 = t[:, 0].min()
    start, blur, copy = (t[:, 0] - t0) / 1e3, (t[:, 1] - t0) / 1e3, (t[:, 2] - t0) / 1e3
    if it >= 3:
        q = lambda a: "min %.1f p10 %.1f med %.1f p90 %.1f max %.1f" % (a.min(), np.percentile(a, 10), np.median(a), np.percentile(a, 90), a.max())
        print("iter %d K1 %.1f us | start: %s | blur end: %s | copy end: %s" % (it, dec.stage_ms()[0] * 1e3, q(start), q(blur), q(copy)))
        if it == 5:
            order = np.argsort(np.maximum(blur, copy))
            print("slowest CTAs (cta, sm, blur_end, copy_end):", [(int(c), int(t[c, 3]), round(float(blur[c]), 1), round(float(copy[c]), 1)) for c in order[-8:]])
            print("fastest CTAs:", [(int(c), int(t[c, 3]), round(float(blur[c]), 1), round(float(copy[c]), 1)) for c in order[:8]])
            # by die / GPC guess: SM id halves
            print("blur end by CTA index (us):", " ".join("%.0f" % v for v in blur))
            print("copy end by CTA index (us):", " ".join("%.0f" % v for v in copy))
            print("SM id by CTA index:", " ".join("%d" % v for v in t[:, 3]))
            for name, sel in (("sm<74", t[:, 3] < 74), ("sm>=74", t[:, 3] >= 74)):
                print(name, "blur end mean %.1f copy end mean %.1f" % (blur[sel].mean(), copy[sel].mean()))
dec.close()
