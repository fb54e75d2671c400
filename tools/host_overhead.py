"""Host time per step of the pipelined decode (three contexts, three streams) in minimal and dense mode: how
long decode_async and fetch take on the host, and the steady-state step time.
usage: python tools/host_overhead.py [steps]"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfpose_b200 as tp
from tfpose_b200 import synth

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 300
h, w, B, H, W = 46, 54, 64, 184, 216
xs = [torch.from_numpy(synth.synth_batch(B, h, w, config_id=2, first_frame=64 * i)).cuda() for i in range(4)]
for prof in (True, False):
    for graph in (0, 1):
        decs = [tp.PoseDecoder(max_batch=B, max_upsample=(H, W)) for _ in range(3)]
        streams = [torch.cuda.Stream() for _ in range(3)]
        for d in decs:
            d.set_profiling(prof)
            d.set_option("k1_ctas", 128)
            try:
                d.set_option("graph", graph)
            except Exception:
                if graph:
                    break
        else:
            for mode in ("minimal", "dense"):
                dense = [tuple(torch.empty((B, H, W, c), device="cuda") for c in (19, 38, 19)) if mode == "dense"
                         else None for _ in range(3)]
                t_dec, t_fetch = [], []
                inflight = []
                torch.cuda.synchronize()
                for i in range(steps + 20):
                    if i == 20:
                        torch.cuda.synchronize()
                        t0 = time.perf_counter()
                    j = i % 3
                    if len(inflight) == 3:
                        a = time.perf_counter()
                        inflight.pop(0).fetch()
                        t_fetch.append(time.perf_counter() - a)
                    a = time.perf_counter()
                    decs[j].decode_async(xs[i % 4], (H, W), dense=dense[j], stream=streams[j])
                    t_dec.append(time.perf_counter() - a)
                    inflight.append(decs[j])
                for d in inflight:
                    d.fetch()
                torch.cuda.synchronize()
                dt = (time.perf_counter() - t0) / steps
                print("profiling %-5s graph %d %-7s step %.1f us (%.0f frames/s)  decode_async %.1f (mean %.1f) us  "
                      "fetch %.1f (mean %.1f) us"
                      % (prof, graph, mode, dt * 1e6, B / dt, np.median(t_dec[20:]) * 1e6, np.mean(t_dec[20:]) * 1e6,
                         np.median(t_fetch[20:]) * 1e6, np.mean(t_fetch[20:]) * 1e6))
                # the GPU side alone: enqueue round-robin without fetching, one synchronize at the end
                for nctx in (1, 2, 3):
                    torch.cuda.synchronize()
                    t0 = time.perf_counter()
                    for i in range(60):
                        decs[i % nctx].decode_async(xs[i % 4], (H, W), dense=dense[i % nctx], stream=streams[i % nctx])
                    t1 = time.perf_counter()
                    torch.cuda.synchronize()
                    t2 = time.perf_counter()
                    print("    enqueue-only over %d ctx/streams: %.1f us/step on the GPU, %.1f us/step to enqueue"
                          % (nctx, (t2 - t0) / 60 * 1e6, (t1 - t0) / 60 * 1e6))
                for d in decs:
                    d.fetch()
                del dense
        for d in decs:
            d.close()
