"""K2 (limb scoring) against its CTAs per frame ("k2_blocks"), sparse (c2) and crowded (c4) frames, 64 frames per call:
K2 timed alone with CUDA events, and the whole pipelined step (dense, three contexts in flight, K1 on 128 CTAs -- the
bench's `value`).  The automatic choice is ~6 CTAs per SM over the batch (14 per frame at batch 64 = 896 CTAs = 1.5
waves of the 592 that fit at four per SM).
usage: python tools/k2_blocks_sweep.py"""
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


def inputs(cfg, people):
    xs = []
    for s in range(4):
        base = synth.synth_batch(8, h, w, config_id=cfg, first_frame=8 * s, **({"people": people} if people else {}))
        xs.append(torch.from_numpy(np.concatenate([base] * 8)).cuda())
    return xs


def alone(xs, caps, bpf):
    dense = tuple(torch.empty((B, H, W, c), device="cuda") for c in (19, 38, 19))
    dec = tp.PoseDecoder(max_batch=B, max_upsample=(H, W), **caps)
    dec.set_profiling(True)
    dec.set_option("k2_blocks", bpf)
    t = []
    for i in range(14):
        dec.decode(xs[i % 4], (H, W), dense=dense)
        if i >= 4:
            t.append(dec.stage_ms()[2])
    dec.close()
    return float(np.mean(t)) * 1e3


def pipelined(xs, caps, bpf, n_pipe=3, steps=60, dense_mode=True):
    decs = [tp.PoseDecoder(max_batch=B, max_upsample=(H, W), **caps) for _ in range(n_pipe)]
    denses = [tuple(torch.empty((B, H, W, c), device="cuda") for c in (19, 38, 19)) for _ in range(n_pipe)] if dense_mode else None
    streams = [torch.cuda.Stream() for _ in range(n_pipe)]
    for d in decs:
        d.set_option("k1_ctas", 128)
        d.set_option("k1_turnstile", 1)
        d.set_option("k2_blocks", bpf)

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
    for name, cfg, people, caps in (("c2", 2, None, {}), ("c4", 4, 30, {"max_peaks": 4096, "max_cands": 8192, "max_humans": 128})):
        xs = inputs(cfg, people)
        for bpf in (0, 4, 6, 8, 9, 10, 12, 14, 18, 24, 32):
            print("%s k2_blocks %2d: K2 alone %6.1f us | pipelined dense step %6.1f us | pipelined minimal (4 ctx) %6.1f us"
                  % (name, bpf, alone(xs, caps, bpf), pipelined(xs, caps, bpf), pipelined(xs, caps, bpf, n_pipe=4, dense_mode=False)), flush=True)
