"""K1 / stage times and achieved GB/s of the fused map kernel for the BASELINE.json config shapes
(dense mode, per-GPU shares): C2 46x54 N=64, C3 46x82 N=64 (128 over 2 GPUs), C4 crowded 46x54 N=64,
C5 92x164 N=64 (512 over 8 GPUs).  usage: python tools/shape_sweep.py [reps]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfpose_b200 as tp
from tfpose_b200 import synth

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 8
cases = [("C2 thin 46x54", 46, 54, 64, 2, None), ("C3 cmu 46x82", 46, 82, 64, 3, None),
         ("C4 crowded 46x54", 46, 54, 64, 4, 30), ("C5 hires 92x164", 92, 164, 64, 5, None)]
for name, h, w, B, cfg, people in cases:
    H, W = 4 * h, 4 * w
    kw = {"people": people} if people else {}
    # four rotating input batches, like bench.py: with the dense outputs of a step they exceed the 126 MB L2, so X is
    # not L2-resident when the map kernel reads it (a single repeated batch made these figures incomparable)
    xs = [torch.from_numpy(np.concatenate([synth.synth_batch(8, h, w, config_id=cfg, first_frame=8 * s_, **kw)] * (B // 8))).cuda()
          for s_ in range(4)]
    dense = (torch.empty((B, H, W, 19), device="cuda"), torch.empty((B, H, W, 38), device="cuda"),
             torch.empty((B, H, W, 19), device="cuda"))
    dec = tp.PoseDecoder(max_batch=B, max_upsample=(H, W), max_peaks=4096, max_cands=8192, max_rows=512, max_humans=128)
    dec.set_profiling(True)
    acc = []
    for i in range(reps + 3):
        res = dec.decode(xs[i % 4], (H, W), dense=dense)
        if i >= 3:
            acc.append(dec.stage_ms())
    st = np.mean(acc, 0)
    alg = 4 * (57 * h * w + 76 * H * W) * B
    dmin = []
    for i in range(reps + 3):
        dec.decode(xs[i % 4], (H, W))
        if i >= 3:
            dmin.append(dec.stage_ms())
    sm = np.mean(dmin, 0)
    print("%-18s N=%d  dense: K1 %.1f us = %.0f GB/s (alg %.0f MB), sort %.1f, K2 %.1f, K3 %.1f | minimal: K1 %.1f us, sort %.1f K2 %.1f K3 %.1f | humans %d peaks/frame %.0f"
          % (name, B, st[0] * 1e3, alg / (st[0] * 1e-3) / 1e9, alg / 1e6, st[1] * 1e3, st[2] * 1e3, st[3] * 1e3,
             sm[0] * 1e3, sm[1] * 1e3, sm[2] * 1e3, sm[3] * 1e3, int(res.info["n_humans"].sum()), res.info["n_peaks"].mean()))
    dec.close()
    del dense, xs
    torch.cuda.empty_cache()
