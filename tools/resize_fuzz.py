"""Random-shape check of the staged strip form of the bicubic resize (the automatic path) against the plain strip form
(path 3: pinned to cv2 through the golden files and the CPU oracle in tests/test_preproc.py), GPU against GPU, bytes
compared.  Shapes: 1-4 channels, source rows 4-byte aligned (what the staged form needs; every fourth case 16-byte
aligned so that its bulk-copy path runs), up- and down-scaling from 1/12 to 6 in either direction, batches of 1-5,
every fifth case through fx / fy instead of a target size.
usage: python tools/resize_fuzz.py [cases] [seed]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfpose_b200 as tp  # noqa: E402


def main():
    cases = int(sys.argv[1]) if len(sys.argv) > 1 else 400
    rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
    d = tp.PoseDecoder(max_batch=1, max_upsample=(184, 216))
    bad = 0
    for i in range(cases):
        cn = int(rng.integers(1, 5))
        step = (16 if i % 4 == 0 else 4) // np.gcd(16 if i % 4 == 0 else 4, cn)
        sw = int(rng.integers(1, 700 // step + 1)) * step
        sh = int(rng.integers(1, 500))
        n = int(rng.integers(1, 6))
        dw = max(1, int(sw * np.exp(rng.uniform(np.log(1 / 12), np.log(6)))))
        dh = max(1, int(sh * np.exp(rng.uniform(np.log(1 / 12), np.log(6)))))
        dw, dh = min(dw, 2000), min(dh, 1500)
        kind = rng.integers(0, 4)
        img = rng.integers(0, 256, (n, sh, sw, cn), dtype=np.uint8)
        if kind == 1:
            img[:, ::2] = 0
            img[:, 1::2] = 255
        elif kind == 2:
            img[:, :, ::2] = 255
            img[:, :, 1::2] = 0
        x = torch.from_numpy(img).cuda()
        out = []
        fx, fy = float(np.float32(dw / sw)) * 1.003, float(np.float32(dh / sh)) * 0.997
        by_factor = i % 5 == 4 and round(sw * fx) >= 1 and round(sh * fy) >= 1
        for path in (0, 3):
            d.set_option("resize_path", path)
            r = d.resize_cubic(x, fx=fx, fy=fy) if by_factor else d.resize_cubic(x, (dw, dh))
            out.append(r.cpu().numpy())
        d.set_option("resize_path", 0)
        if not np.array_equal(out[0], out[1]):
            bad += 1
            diff = np.argwhere(out[0] != out[1])
            print("MISMATCH case %d: n=%d %dx%dx%d -> %dx%d: %d bytes differ, first at %s (%d vs %d)"
                  % (i, n, sh, sw, cn, dh, dw, len(diff), diff[0].tolist(), out[0][tuple(diff[0])], out[1][tuple(diff[0])]))
    d.close()
    print("resize fuzz: %d cases, %d mismatches" % (cases, bad))
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
