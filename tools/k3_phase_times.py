"""Per-phase cycle counts of the assembly kernel (TFPD_K1_TIMES probe buffer, K3 part).
usage: python tools/k3_phase_times.py [config=2|4]"""
import os
import sys

import numpy as np
import torch

os.environ.setdefault("TFPD_K1_TIMES", "1")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfpose_b200 as tp
from tfpose_b200 import synth

CFG = int(sys.argv[1]) if len(sys.argv) > 1 else 2
B, h, w = 64, 46, 54
H, W = 4 * h, 4 * w
base = synth.synth_batch(16, h, w, config_id=CFG, **({"people": 30} if CFG == 4 else {}))
x = torch.from_numpy(np.concatenate([base] * 4)).cuda()
dec = tp.PoseDecoder(max_batch=B, max_upsample=(H, W), max_peaks=4096, max_cands=8192, max_humans=128)
dec.set_profiling(True)
names = ["load", "bucket", "sort", "greedy", "assembly", "filter+out"]
for it in range(5):
    res = dec.decode(x, (H, W))
    t = dec.k1_cta_times(1024).astype(np.int64).reshape(-1)[2048:2048 + 64 * 8].reshape(64, 8)
    d = np.diff(t[:, :7], axis=1)
    if it >= 3:
        tot = t[:, 6] - t[:, 0]
        slow = int(np.argmax(tot))
        print("iter %d K3 %.1f us; cycles mean per phase: %s | slowest frame %d (%d rows, %d cands): %s total %d"
              % (it, dec.stage_ms()[3] * 1e3, dict(zip(names, d.mean(0).astype(int).tolist())), slow,
                 int(res.info["n_rows"][slow]), int(res.info["n_cands"][slow]), dict(zip(names, d[slow].tolist())), tot[slow]))
        print("   slowest frame: %d connections, longest limb segment %d candidates" % (t[slow, 7] >> 32, t[slow, 7] & 0xffffffff))
dec.close()
