"""Sweeps on the bench workload (c2, 64 frames, dense, rotating inputs): K1 time against its CTA count, and the
pipelined step time against the number of contexts in flight and the CTA count (c2 and the crowded c4).
usage: python tools/k1_sweep.py            everything
       python tools/k1_sweep.py c4         only the crowded pipelined lines (run once with TFPD_K3_MINB=1 and once
                                           without to compare K3 at one / two CTAs per SM: the variant is chosen at
                                           process start)"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfpose_b200 as tp
from tfpose_b200 import synth

B, h, w = 64, 46, 54
H, W = 4 * h, 4 * w


def inputs(cfg, people=None):
    xs = []
    for s in range(4):
        base = synth.synth_batch(8, h, w, config_id=cfg, first_frame=8 * s, **({"people": people} if people else {}))
        xs.append(torch.from_numpy(np.concatenate([base] * 8)).cuda())
    return xs


def k1_alone(xs):
    dense = tuple(torch.empty((B, H, W, c), device="cuda") for c in (19, 38, 19))
    dec = tp.PoseDecoder(max_batch=B, max_upsample=(H, W))
    dec.set_profiling(True)
    for ctas in (96, 104, 112, 120, 124, 128, 132, 136, 140, 148):
        dec.set_option("k1_ctas", ctas)
        t = []
        for i in range(15):
            dec.decode(xs[i % 4], (H, W), dense=dense)
            if i >= 3:
                t.append(dec.stage_ms()[0])
        print("K1 dense alone, %3d CTAs: %.1f us (min %.1f)" % (ctas, np.mean(t) * 1e3, np.min(t) * 1e3))
    dec.close()


def pipelined(xs, n_pipe, ctas, steps=60, dense_mode=True, tail_priority=0, band=4):
    decs = [tp.PoseDecoder(max_batch=B, max_upsample=(H, W)) for _ in range(n_pipe)]
    denses = [tuple(torch.empty((B, H, W, c), device="cuda") for c in (19, 38, 19)) for _ in range(n_pipe)] if dense_mode else None
    streams = [torch.cuda.Stream() for _ in range(n_pipe)]
    for d in decs:
        d.set_option("k1_ctas", ctas)
        d.set_option("k1_turnstile", 1)
        d.set_option("tail_priority", tail_priority)
        d.set_option("k1_min_band", band)

    def run(n):
        inflight = []
        for i in range(n):
            j = i % n_pipe
            if len(inflight) == n_pipe:
                inflight.pop(0).fetch(reuse=True)
            decs[j].decode_async(xs[i % 4], (H, W), dense=denses[j] if dense_mode else None, stream=streams[j])
            inflight.append(decs[j])
        for d in inflight:
            d.fetch(reuse=True)
    run(2 * n_pipe)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    run(steps)
    torch.cuda.synchronize()
    us = (time.perf_counter() - t0) / steps * 1e6
    for d in decs:
        d.close()
    return us


if __name__ == "__main__":
    only_c4 = len(sys.argv) > 1 and sys.argv[1] == "c4"
    x2, x4 = inputs(2), inputs(4, 30)
    if not only_c4:
        k1_alone(x2)
    print("TFPD_K3_MINB =", os.environ.get("TFPD_K3_MINB", "(automatic: two CTAs per SM when the shared memory fits)"))
    for name, xs in ((("c4", x4),) if only_c4 else (("c2", x2), ("c4", x4))):
        for ctas in (112, 120, 128, 148):
            print("%s pipelined dense, %3d CTAs: " % (name, ctas) + "  ".join(
                "%d ctx %.1f us" % (n, pipelined(xs, n, ctas)) for n in (2, 3, 4, 6)))
        for tp_ in (0, 1):
            for band in (4, 8):
                print("%s pipelined minimal, tail_priority %d, band %d: " % (name, tp_, band) + "  ".join(
                    "%d ctx %.1f us" % (n, pipelined(xs, n, 0, dense_mode=False, tail_priority=tp_, band=band)) for n in (2, 3, 4, 6)))
        print("%s pipelined dense, 128 CTAs, tail_priority 1: " % name + "  ".join(
            "%d ctx %.1f us" % (n, pipelined(xs, n, 128, tail_priority=1)) for n in (2, 3, 4)))
