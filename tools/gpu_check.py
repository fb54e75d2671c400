"""Developer diagnostic (not a test): decode a few synthetic frames on the GPU and print how each
stage compares with the oracle.  Run on the GPU box: python tools/gpu_check.py"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfpose_b200 as tp
from tfpose_b200 import synth
from oracle import oracle as O


def check(N, h, w, people, dense_mode, cfgid=2):
    H, W = 4 * h, 4 * w
    x = synth.synth_batch(N, h, w, config_id=cfgid, people=people)
    dec = tp.PoseDecoder(max_batch=N, max_upsample=(H, W), device=0, max_peaks=4096, max_cands=8192, max_humans=128)
    dec.set_profiling(True)
    xd = torch.from_numpy(x).cuda()
    dense = None
    if dense_mode:
        dense = (torch.full((N, H, W, 19), 7.0, device="cuda"), torch.full((N, H, W, 38), 7.0, device="cuda"),
                 torch.full((N, H, W, 19), 7.0, device="cuda"))
    t = time.time()
    res = dec.decode(xd, (H, W), dense=dense)
    torch.cuda.synchronize()
    print("N=%d %dx%d people=%s dense=%s: %.2f ms wall, stages %s, launches %d" % (
        N, h, w, people, dense_mode, (time.time() - t) * 1e3, ["%.3f" % v for v in dec.stage_ms()], dec.launch_count()))
    bad = 0
    for i in range(min(N, 3)):
        R, (peaks, heat_up, paf_up, g) = O.decode_frame(x[i], H, W, sort="total")
        pk = dec.peaks(i)
        print("  frame %d: info %s" % (i, res.info[i]))
        same_peaks = (len(pk["x"]) == len(R.peaks["x"]) and np.array_equal(pk["x"], R.peaks["x"]) and
                      np.array_equal(pk["y"], R.peaks["y"]) and np.array_equal(pk["part"], R.peaks["part"]) and
                      np.array_equal(pk["score"], R.peaks["score"]))
        print("    peaks gpu %d oracle %d identical=%s fragile=%d" % (len(pk["x"]), len(R.peaks["x"]), same_peaks,
                                                                   int(O.fragile_pixels(g).sum())))
        if dense_mode:
            hu, pu, pkd = (t_[i].cpu().numpy() for t_ in dense)
            print("    heat_up exact=%s paf_up exact=%s peaks max err=%.3e (rel %.3e) nonzero gpu/oracle %d/%d" % (
                np.array_equal(hu, heat_up), np.array_equal(pu, paf_up), np.abs(pkd - peaks).max(),
                np.abs(pkd - peaks).max() / np.abs(g).max(), int((pkd != 0).sum()), int((peaks != 0).sum())))
            mism = np.argwhere((pkd != 0) != (peaks != 0))
            if len(mism):
                print("    peak-support mismatches:", len(mism), mism[:5].tolist())
        conns = dec.connections(i)
        oc = [(l, int(a), int(b), float(s)) for l in range(19) for a, b, s in zip(*R.conns[l])]
        gc = list(zip(conns["limb"].tolist(), conns["cid1"].tolist(), conns["cid2"].tolist(), conns["score"].tolist()))
        print("    conns gpu %d oracle %d identical=%s" % (len(gc), len(oc), gc == [(l, a, b, np.float32(s)) for l, a, b, s in oc]))
        got, want = res.humans(i), R.humans_as_tuples()
        ok = len(got) == len(want)
        if ok:
            for hg, (score, parts) in zip(got, want):
                ok &= np.float32(hg.score) == np.float32(score) and sorted(hg.body_parts) == sorted(parts)
                if ok:
                    for k, (px, py, ps) in parts.items():
                        bp = hg.body_parts[k]
                        ok &= bp.x == px / W and bp.y == py / H and np.float32(bp.score) == np.float32(ps)
        print("    humans gpu %d oracle %d identical=%s" % (len(got), len(want), ok))
        bad += 0 if ok else 1
    dec.close()
    return bad


if __name__ == "__main__":
    bad = 0
    bad += check(2, 46, 54, None, True)
    bad += check(2, 46, 54, None, False)
    bad += check(3, 46, 82, 5, True, cfgid=3)
    bad += check(2, 46, 54, 30, True, cfgid=4)
    bad += check(2, 92, 164, 30, True, cfgid=5)
    bad += check(64, 46, 54, None, True)
    print("FAILED FRAMES:", bad)
