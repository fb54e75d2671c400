"""Stage times of the general resize path (k_generic.cu) next to the x4 fast path: 46x54x57 maps, upsample
ratios 8 (eval.py's default), 2 and 3, dense and minimal mode.  usage: python tools/generic_ratio_times.py [ratio ...]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfpose_b200 as tp
from tfpose_b200 import synth

h, w, B = 46, 54, 16
x = torch.from_numpy(synth.synth_batch(B, h, w, config_id=2)).cuda()
ratios = [int(a) for a in sys.argv[1:]] or [4, 8, 2, 16, 1, 3]
for ratio in ratios:
    H, W = h * ratio, w * ratio
    dec = tp.PoseDecoder(max_batch=B, max_upsample=(H, W), max_peaks=4096, max_cands=8192)
    dec.set_profiling(True)
    if os.environ.get("POW2_PATH"):
        dec.set_option("pow2_path", int(os.environ["POW2_PATH"]))
    dense = tuple(torch.empty((B, H, W, c), device="cuda") for c in (19, 38, 19))
    for mode, d in (("dense", dense), ("minimal", None)):
        acc = []
        for i in range(6):
            res = dec.decode(x, (H, W), dense=d)
            if i >= 2:
                acc.append(dec.stage_ms())
        st = np.mean(acc, 0) * 1e3
        alg = 4 * (57 * h * w + 76 * H * W) * B
        print("ratio %d %-7s maps %.1f us (%s), sort %.1f, K2 %.1f, K3 %.1f, humans %d, launches %d"
              % (ratio, mode, st[0], ("%.0f GB/s" % (alg / st[0] / 1e3)) if d else "-", st[1], st[2], st[3],
                 int(res.info["n_humans"].sum()), dec.launch_count()))
    dec.close()
    del dense
