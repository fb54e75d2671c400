"""Times the pre-processing kernels (k_resize_cubic_u8, k_quantize_img) on device tensors with CUDA events.
usage: python tools/preproc_times.py [batch] [src_h src_w]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfpose_b200 as tp  # noqa: E402


def timed(fn, reps=50, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    sh, sw = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (480, 640)
    d = tp.PoseDecoder(max_batch=1, max_upsample=(184, 216))
    # 8 rotating batches of source images: 8 x n x 0.92 MB, beyond the 126 MB L2 for n = 64
    srcs = [torch.randint(0, 256, (n, sh, sw, 3), dtype=torch.uint8, device="cuda") for _ in range(8)]
    i = [0]

    def resize():
        i[0] += 1
        return d.resize_cubic(srcs[i[0] % 8], (432, 368))

    by = n * (sh * sw * 3 + 368 * 432 * 3)
    for path, name in ((0, "staged strip (automatic)"), (3, "strip"), (1, "tiled"), (2, "per-pixel")):
        d.set_option("resize_path", path)
        ms = timed(resize)
        print("resize_cubic %s %dx%dx3 -> 368x432x3, batch %d: %.3f ms, %.0f frames/s, %.0f GB/s algorithmic" %
              (name, sh, sw, n, ms, n / ms * 1e3, by / ms / 1e6))
    d.set_option("resize_path", 0)
    xs = [torch.rand((n, 368, 432, 3), device="cuda") * 2 - 1 for _ in range(4)]

    def quant():
        i[0] += 1
        return d.quantize_img(xs[i[0] % 4])

    ms = timed(quant)
    print("quantize_img batch %d: %.3f ms, %.0f GB/s" % (n, ms, n * 368 * 432 * 3 * 5 / ms / 1e6))
    d.close()
    try:                                            # the reference's own call on this box's host cores
        import time

        import cv2
        imgs = srcs[0].cpu().numpy()
        for ipp in (True, False):
            cv2.ipp.setUseIPP(ipp)
            t = time.perf_counter()
            for im in imgs:
                cv2.resize(im, (432, 368), interpolation=cv2.INTER_CUBIC)
            dt = time.perf_counter() - t
            print("cv2.resize INTER_CUBIC on the host (%d threads, IPP %s): %.3f ms per frame, %.0f frames/s" %
                  (cv2.getNumThreads(), "on" if ipp else "off", dt / n * 1e3, n / dt))
    except ImportError:
        pass


if __name__ == "__main__":
    main()
