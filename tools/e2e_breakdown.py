"""Where the time of one decode_host call goes (bench workload, pinned input)."""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfpose_b200 as tp
from tfpose_b200 import synth, _lib

B, h, w = 64, 46, 54
H, W = 4 * h, 4 * w
base = synth.synth_batch(16, h, w, config_id=2)
xp = torch.from_numpy(np.concatenate([base] * 4)).pin_memory()
x = xp.numpy()
xd = torch.empty_like(xp, device="cuda")
dec = tp.PoseDecoder(max_batch=B, max_upsample=(H, W))


def timeit(fn, n=30, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3


print("pure H2D 36.2 MB pinned        %.3f ms" % timeit(lambda: xd.copy_(xp, non_blocking=True)))
print("decode_host (python)           %.3f ms" % timeit(lambda: dec.decode_host(x, (H, W))))
L = dec._L
raw = lambda: L.tfpd_decode_host(dec._ctx, x.ctypes.data, B, h, w, H, W, None, None, None)
print("tfpd_decode_host (raw ctypes)  %.3f ms" % timeit(raw))
raw()
dec._shape = (B, H, W)
print("fetch() after the call         %.3f ms" % timeit(lambda: dec.fetch()))
print("decode (device input, minimal) %.3f ms" % timeit(lambda: dec.decode(xd, (H, W))))
dec.set_profiling(True)
dec.decode(xd, (H, W))
print("stage ms minimal mode:", dec.stage_ms())
for n in (8, 16, 32):
    dec.decode(xd[:n], (H, W))
    print("stage ms minimal mode, %d frames:" % n, ["%.3f" % v for v in dec.stage_ms()])
dec.close()

# ---- python-side pieces of decode_host ----
dec = tp.PoseDecoder(max_batch=B, max_upsample=(H, W))
import cProfile, pstats
for _ in range(5):
    dec.decode_host(x, (H, W))
pr = cProfile.Profile()
pr.enable()
for _ in range(30):
    dec.decode_host(x, (H, W))
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(14)
t = []
for _ in range(30):
    t0 = time.perf_counter(); a = np.ascontiguousarray(x, np.float32); t1 = time.perf_counter()
    p = a.ctypes.data; t2 = time.perf_counter()
    info = np.empty(B, _lib.INFO_DTYPE); table = np.empty((B, dec.max_humans), _lib.HUMAN_DTYPE); t3 = time.perf_counter()
    t.append((t1 - t0, t2 - t1, t3 - t2))
print("ascontiguousarray %.1f us, ctypes.data %.1f us, np.empty x2 %.1f us" % tuple(np.mean(t, 0) * 1e6))
for ch in (1, 4):
    os.environ["TFPD_HOST_CHUNKS"] = str(ch)
    d2 = tp.PoseDecoder(max_batch=B, max_upsample=(H, W))
    f = lambda: d2._L.tfpd_decode_host(d2._ctx, x.ctypes.data, B, h, w, H, W, None, None, None)
    print("raw tfpd_decode_host, %d chunk(s): %.3f ms" % (ch, timeit(f)))
    d2.close()
