"""CPU check of the exactness argument behind K1's minimal-mode early-out ("k1_early_out", include/tfpose_decode.h,
bound derived in csrc/api.cu): on seeded synthetic frames of the BASELINE shapes, every (channel, low-res row group,
column chunk) unit the kernel's rule would skip -- largest heat value in the 8 low-res rows the group reads, over the
cells the warp holds, <= thresh * (1 - 2e-4) / sum_w^2 -- is checked against the ORACLE's blurred map: no blurred
value of the five output rows the group computes or decides may come within 2e-5 relative of thresh * (1 - 4e-5).
Also evaluates a tighter rule (row maxima weighted by the vertical smoother weights) that was considered and NOT
adopted: it is exact as well but skips only ~6 points more.  Test infrastructure (uses oracle/); no GPU needed.
usage: python tools/check_early_out_bound.py"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tfpose_b200 import synth
from oracle import oracle as O
_, s1d = O.gauss_kernel()
wp = np.zeros((4, 7), np.float32)
for p in range(4):
    acc = np.zeros(7)
    for q in range(25):
        acc[(p + q - 12) // 4 + 3] += s1d[q]
    wp[p] = acc.astype(np.float32)
sw = max(float(wp[p].astype(np.float64).sum()) for p in range(4))
thr = np.float32(0.05); thr_lo = thr * np.float32(1 - 4e-5)
limit = float(thr) * (1 - 2e-4) / sw            # bound on sum_t wv[t] * rowmax_t  (one factor sw for the horizontal sum)
simple_bound = float(thr) * (1 - 2e-4) / (sw * sw)
stats = {}
for name, frames in (("c2", [(46, 54, 2, None, i) for i in range(12)]), ("c4", [(46, 54, 4, 30, i) for i in range(4)]),
                     ("c3", [(46, 82, 3, None, i) for i in range(4)])):
    tot = d1 = d2 = 0
    worst = worst1 = 0.0
    for (h, w, cfg, people, fr) in frames:
        x = synth.synth_frame(h, w, synth.frame_seed(cfg, fr), people=people)
        H, W = 4 * h, 4 * w
        _, _, _, g = O.dense_maps(x, H, W)
        hp = np.zeros((h + 16, w, 19), np.float32); hp[8:8 + h] = x[:, :, :19]
        nch = (w + 55) // 56; cc = (w + nch - 1) // nch
        for iy in range(h):
            for ch in range(nch):
                c0, c1 = ch * cc, min(w, (ch + 1) * cc)
                lo, hi = max(0, c0 - 4), min(w, c1 + 4)
                rows = hp[8 + iy - 4: 8 + iy + 4, lo:hi, :]            # low-res rows iy-4 .. iy+3
                m = np.maximum(rows.max(axis=1), 0.0).astype(np.float64)   # [8, 19] row maxima over the held cells, clamped at 0
                # rows 4iy-2, 4iy-1: phases 2,3 around centre iy-1 (rows iy-4..iy+2 = m[0:7]); rows 4iy..4iy+2: phases 0,1,2 around iy (m[1:8])
                b = np.zeros(19)
                for p in (2, 3):
                    b = np.maximum(b, (wp[p].astype(np.float64)[:, None] * m[0:7]).sum(0))
                for p in (0, 1, 2):
                    b = np.maximum(b, (wp[p].astype(np.float64)[:, None] * m[1:8]).sum(0))
                dead2 = b <= limit
                dead1 = rows.max(axis=(0, 1)) <= simple_bound
                assert (dead2 | ~dead1).all()                          # the tighter rule skips a superset
                y0, y1 = max(0, 4 * iy - 2), min(H, 4 * iy + 3)
                gg = g[y0:y1, 4 * c0:4 * c1, :].max(axis=(0, 1))
                tot += 19; d1 += int(dead1.sum()); d2 += int(dead2.sum())
                if dead1.any():
                    worst1 = max(worst1, float((gg[dead1] / thr_lo).max()))
                if dead2.any():
                    worst = max(worst, float((gg[dead2] / thr_lo).max()))
                    assert (gg[dead2] < thr_lo * np.float32(1 - 2e-5)).all(), (cfg, fr, iy, ch)
    print("%s: units %d | ADOPTED rule (window max): skipped %.1f %%, largest blurred value in a skipped unit %.3f x thr_lo | "
          "row-weighted rule (not adopted): skipped %.1f %%, largest %.3f x thr_lo"
          % (name, tot, 100.0 * d1 / tot, worst1, 100.0 * d2 / tot, worst))
