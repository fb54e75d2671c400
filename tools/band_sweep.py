"""Minimal mode: whole-step time against the band size of the map kernel ("k1_min_band"), for every BASELINE shape and
several batch sizes, one context (serial) and four contexts in flight (the bench's pipelined figure).

K1 ALONE prefers ~4 waves of small CTAs (tools/k1_ab.py: c2 50.0 us at band 4, 51.5 at band 8); with other batches'
K2 / K3 competing for the SMs what counts is K1's total SM time, which falls with the band size (fewer halo rows and
window prologues).  This sweep is what the automatic choice in api.cu is derived from.
usage: python tools/band_sweep.py"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfpose_b200 as tp
from tfpose_b200 import synth

SHAPES = (("c2", 46, 54, 2, None, {}), ("c4", 46, 54, 4, 30, {"max_peaks": 4096, "max_cands": 8192, "max_humans": 128}),
          ("c3", 46, 82, 3, None, {}), ("c5", 92, 164, 5, None, {}))


def inputs(B, h, w, cfg, people):
    xs = []
    per = min(8, B)
    for s in range(4):
        base = synth.synth_batch(per, h, w, config_id=cfg, first_frame=per * s, **({"people": people} if people else {}))
        xs.append(torch.from_numpy(np.concatenate([base] * (B // per))).cuda())
    return xs


def step_us(xs, B, H, W, caps, n_pipe, band, steps=60):
    decs = [tp.PoseDecoder(max_batch=B, max_upsample=(H, W), **caps) for _ in range(n_pipe)]
    streams = [torch.cuda.Stream() for _ in range(n_pipe)]
    for d in decs:
        d.set_option("k1_min_band", band)

    def run(n):
        inflight = []
        for i in range(n):
            j = i % n_pipe
            if len(inflight) == n_pipe:
                inflight.pop(0).fetch(reuse=True)
            decs[j].decode_async(xs[i % 4], (H, W), stream=streams[j])
            inflight.append(decs[j])
        for d in inflight:
            d.fetch(reuse=True)
    run(2 * n_pipe + 2)
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(2):
        t0 = time.perf_counter()
        run(steps)
        torch.cuda.synchronize()
        best = min(best, (time.perf_counter() - t0) / steps * 1e6)
    for d in decs:
        d.close()
    return best


if __name__ == "__main__":
    for name, h, w, cfg, people, caps in SHAPES:
        for B in (64, 16, 8):
            xs = inputs(B, h, w, cfg, people)
            for n_pipe in (1, 4):
                row = []
                for band in (0, 2, 4, 8, 16, 32):
                    row.append("band %2d %6.1f" % (band, step_us(xs, B, 4 * h, 4 * w, caps, n_pipe, band)))
                print("%s B=%2d %d ctx (us per step; band 0 = automatic): %s" % (name, B, n_pipe, "  ".join(row)), flush=True)
            del xs
            torch.cuda.empty_cache()
