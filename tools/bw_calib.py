"""Calibration: what a plain fill (write-only) and a plain copy (read+write) reach on this GPU."""
import torch
n = 1 << 30
a = torch.empty(n, dtype=torch.uint8, device="cuda")
b = torch.empty(n, dtype=torch.uint8, device="cuda")
def timeit(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best
t = timeit(lambda: a.zero_()); print("zero_ 1GiB: %.3f ms -> %.0f GB/s write-only" % (t, n / t / 1e6))
t = timeit(lambda: a.fill_(7)); print("fill_ 1GiB: %.3f ms -> %.0f GB/s write-only" % (t, n / t / 1e6))
t = timeit(lambda: b.copy_(a)); print("copy_ 1GiB: %.3f ms -> %.0f GB/s read+write" % (t, 2 * n / t / 1e6))
m = 773 * 1000 * 1000
c = torch.empty(m, dtype=torch.uint8, device="cuda")
t = timeit(lambda: c.zero_()); print("zero_ 773MB: %.3f ms -> %.0f GB/s write-only" % (t, m / t / 1e6))
af = a.view(torch.float32)
t = timeit(lambda: af.sum()); print("sum 1GiB f32: %.3f ms -> %.0f GB/s read-only" % (t, n / t / 1e6))
