"""Small profiling target: the bench workload (64 mobilenet_thin frames, dense mode), W warm-up
and K profiled decode steps, nothing else.  Usage: python tools/ncu_target.py [K] [W] [dense=1] [config=2]
(config 4 = crowded frames, ~30 people: the limb-scoring / assembly stress case of BASELINE.json)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfpose_b200 as tp
from tfpose_b200 import synth
import numpy as np

K = int(sys.argv[1]) if len(sys.argv) > 1 else 3
Wm = int(sys.argv[2]) if len(sys.argv) > 2 else 3
DENSE = int(sys.argv[3]) if len(sys.argv) > 3 else 1
CFG = int(sys.argv[4]) if len(sys.argv) > 4 else 2
B, h, w = 64, 46, 54
H, W = 4 * h, 4 * w
base = synth.synth_batch(16, h, w, config_id=CFG, **({"people": 30} if CFG == 4 else {}))
x = torch.from_numpy(np.concatenate([base] * 4)).cuda()
dense = None
if DENSE:
    dense = (torch.empty((B, H, W, 19), device="cuda"), torch.empty((B, H, W, 38), device="cuda"),
             torch.empty((B, H, W, 19), device="cuda"))
dec = tp.PoseDecoder(max_batch=B, max_upsample=(H, W), **({"max_peaks": 4096, "max_cands": 8192, "max_humans": 128} if CFG == 4 else {}))
dec.set_profiling(True)
dec.set_option("k1_ctas", int(os.environ.get("K1_CTAS", "128")))      # the bench setting
for i in range(Wm + K):
    res = dec.decode(x, (H, W), dense=dense)
    print("step", i, ["%.3f" % v for v in dec.stage_ms()], int(res.info["n_humans"].sum()))
dec.close()
