"""Small workload for compute-sanitizer (tools/sanitize_gpu.sh): one pass over every kernel path of the library
at sizes a sanitizer finishes in minutes, each result checked against the CPU oracle so that a run that "passes"
under the tool also computed the right thing.

  K1 fused map kernel   dense and minimal mode (early-out on and off), a ragged width (two column chunks, partial
                        copy items), a batch that spans several CTAs' row spans
  K2 limb scoring       fused peak ranking and the separate sort kernel, sparse and crowded frames
  K3 assembly           as-built and upstream mode, crowded frame (> 16 candidates per limb, found == 2 merges),
                        register rows and the shared-memory fall-back (TFPD_K3_REGROWS=0 in a second process)
  general path          power-of-two ratios 2 and 8, the generic ratio-3 kernels
  seam 2                process_paf on dense host arrays (k_scan_dense_peaks)
  next rows             draw_humans, bicubic resize (strip form), quantize
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfpose_b200 as tp
from tfpose_b200 import synth
from oracle import oracle as O

CAPS = dict(max_peaks=4096, max_cands=8192, max_humans=128)


def same_humans(res, i, want, H, W):
    got = res.humans(i)
    assert len(got) == len(want), (len(got), len(want))
    for hg, (score, parts) in zip(got, want):
        assert np.float32(hg.score) == np.float32(score) and sorted(hg.body_parts) == sorted(parts)
        for k, (px, py, ps) in parts.items():
            bp = hg.body_parts[k]
            assert bp.x == px / W and bp.y == py / H and np.float32(bp.score) == np.float32(ps)


def decode_and_check(x, factor, mode="asbuilt", dense_mode=True, options=(), **cfg):
    n, h, w, _ = x.shape
    H, W = factor * h, factor * w
    dec = tp.PoseDecoder(max_batch=n, max_upsample=(H, W), mode=mode, **cfg)
    for k, v in options:
        dec.set_option(k, v)
    xd = torch.from_numpy(x).cuda()
    dense = tuple(torch.empty((n, H, W, c), device="cuda") for c in (19, 38, 19)) if dense_mode else None
    res = dec.decode(xd, (H, W), dense=dense)
    torch.cuda.synchronize()
    checked = 0
    for i in range(n):
        R, (peaks, hu, pu, g) = O.decode_frame(x[i], H, W, mode=mode, sort="std")
        if dense_mode:
            assert np.array_equal(dense[0][i].cpu().numpy(), hu) and np.array_equal(dense[1][i].cpu().numpy(), pu)
        if int(O.fragile_pixels(g).sum()) == 0 and not (R.flags & 1):
            same_humans(res, i, R.humans_as_tuples(), H, W)
            checked += 1
    dec.close()
    return checked


def main():
    total = 0
    x2 = synth.synth_batch(3, 46, 54, config_id=2, first_frame=2)
    for dense_mode in (True, False):
        total += decode_and_check(x2, 4, dense_mode=dense_mode)
    total += decode_and_check(x2, 4, dense_mode=False, options=(("k1_early_out", 0),))
    total += decode_and_check(x2, 4, options=(("fuse_sort", 0),))
    total += decode_and_check(synth.synth_batch(2, 46, 82, config_id=3, first_frame=1), 4)           # two column chunks
    rng = np.random.default_rng(5113)
    xr = rng.random((2, 5, 113, 57), dtype=np.float32)
    xr[..., 19:] = xr[..., 19:] * 2 - 1
    total += decode_and_check(xr, 4, max_peaks=4096, max_cands=8192, max_rows=1024, max_humans=512)  # ragged, every pixel hot
    xc = synth.synth_batch(1, 46, 54, config_id=4, people=30, first_frame=3)
    for mode in ("asbuilt", "upstream"):
        for dense_mode in (True, False):
            total += decode_and_check(xc, 4, mode=mode, dense_mode=dense_mode, **CAPS)
    xs = synth.synth_batch(1, 23, 27, config_id=2, first_frame=9)
    total += decode_and_check(xs, 2)
    total += decode_and_check(xs, 8)
    total += decode_and_check(xs, 3)
    print("decode paths: %d frames matched the oracle exactly" % total)

    # seam 2: dense host arrays through tfpd_process_paf (k_scan_dense_peaks)
    from tfpose_b200 import pafprocess as pp
    pk, hu, pu, g = O.dense_maps(x2[0], 184, 216)
    humans = tp.PoseEstimator.estimate_paf(pk, hu, pu)
    R = O.process_paf(pk, hu, pu, sort="std")
    assert len(humans) == len(R.humans_as_tuples())
    pp.configure()
    print("seam 2: %d humans" % len(humans))

    # draw_humans, bicubic resize, quantize
    dec = tp.PoseDecoder(max_batch=2, max_upsample=(184, 216))
    img = np.zeros((120, 160, 3), np.uint8)
    dec.draw_humans(img, humans)
    assert img.any() or not humans
    src = rng.integers(0, 256, (2, 97, 131, 3), dtype=np.uint8)
    from oracle import preproc_oracle as PO
    out = dec.resize_cubic(src, (54, 46))
    assert np.array_equal(out[0], PO.resize_cubic_u8(src[0], 54, 46))
    q = dec.quantize_img(rng.random((3, 17, 19), dtype=np.float32) * 2 - 1)
    assert q.dtype == np.uint8
    dec.close()
    print("sanitize target OK")


if __name__ == "__main__":
    main()
