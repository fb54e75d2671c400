"""Stage times on crowded frames (C4: ~30 people per frame), with and without the fused peak ranking."""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfpose_b200 as tp
from tfpose_b200 import synth
B, h, w = 64, 46, 54
x = torch.from_numpy(synth.synth_batch(B, h, w, config_id=4, people=30)).cuda()
dense = tuple(torch.empty((B, 184, 216, c), device="cuda") for c in (19, 38, 19))
for fuse in (0, 1):
    dec = tp.PoseDecoder(max_batch=B, max_upsample=(184, 216), max_peaks=4096, max_cands=16384)
    dec.set_profiling(True)
    dec.set_option("fuse_sort", fuse)
    acc = []
    for i in range(8):
        res = dec.decode(x, (184, 216), dense=dense)
        if i >= 3:
            acc.append(dec.stage_ms())
    st = np.mean(acc, 0) * 1e3
    print("crowded fuse_sort %d: K1 %.1f sort %.1f K2 %.1f K3 %.1f | peaks/frame %.0f humans %d" % (fuse, st[0], st[1], st[2], st[3], res.info["n_peaks"].mean(), res.info["n_humans"].sum()))
    dec.close()
