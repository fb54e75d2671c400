"""draw_humans (tf_pose/estimator.py:391-415) as CUDA kernels on a headless buffer, SURVEY.md 8f rank 3.

Checker = OpenCV itself (the reference delegates the rasterisation to cv2.circle / cv2.line; cv2 is part of the
image): `reference_draw` below is the reference's loop, statement for statement, over cv2.  The CPU test builds
csrc/draw_raster.h (the integer rasterisation rules the kernel runs) with g++ and compares them with cv2 on random
primitives, including clipped ones; the GPU tests compare whole drawn images bit for bit."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import tfpose_b200 as tp
from tfpose_b200 import common

cv2 = pytest.importorskip("cv2")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def reference_draw(npimg, humans):
    """TfPoseEstimator.draw_humans, estimator.py:394-415."""
    image_h, image_w = npimg.shape[:2]
    centers = {}
    for human in humans:
        for i in range(common.CocoPart.Background.value):
            if i not in human.body_parts.keys():
                continue
            body_part = human.body_parts[i]
            center = (int(body_part.x * image_w + 0.5), int(body_part.y * image_h + 0.5))
            centers[i] = center
            cv2.circle(npimg, center, 3, common.CocoColors[i], thickness=3, lineType=8, shift=0)
        for pair_order, pair in enumerate(common.CocoPairsRender):
            if pair[0] not in human.body_parts.keys() or pair[1] not in human.body_parts.keys():
                continue
            cv2.line(npimg, centers[pair[0]], centers[pair[1]], common.CocoColors[pair_order], 3)
    return npimg


def random_humans(rng, n, spread=1.0):
    from tfpose_b200 import synth
    out = []
    for i in range(n):
        hm = tp.Human([])
        keep = rng.random(18) < rng.uniform(0.4, 1.0)
        cx, cy, size = rng.uniform(-0.05, 1.05), rng.uniform(-0.05, 0.8), rng.uniform(0.1, 0.9) * spread
        for k in range(18):
            if keep[k]:
                px = float(np.clip(cx + synth._TEMPLATE[k, 0] * size, 0.0, 0.999999))
                py = float(np.clip(cy + synth._TEMPLATE[k, 1] * size, 0.0, 0.999999))
                hm.body_parts[k] = tp.BodyPart('%d-%d' % (i, k), k, px, py, 0.9)
        out.append(hm)
    return out


def test_raster_rules_match_cv2(tmp_path):
    """csrc/draw_raster.h compiled for the host == cv2.line(thickness 3) / cv2.circle(r 3, thickness 3)."""
    src = tmp_path / "host.cpp"
    src.write_text("""#include "draw_raster.h"
using namespace tfpd_draw;
struct Put { unsigned char *img; int W; void operator()(int x, int y) { img[y * W + x] = 255; } };
extern "C" void h_line(unsigned char *img, int W, int H, int x0, int y0, int x1, int y1)
{ Put p{img, W}; cv_line(W, H, x0, y0, x1, y1, 3, p); }
extern "C" void h_circle3(unsigned char *img, int W, int H, int cx, int cy)
{ Put p{img, W}; for (int s = 0; s < 12; ++s) cv_circle3_segment(W, H, cx, cy, s, p); }
""")
    lib = tmp_path / "libh.so"
    subprocess.run(["g++", "-O2", "-fPIC", "-shared", "-std=c++17", "-I",
                    os.path.join(ROOT, "tf-pose-estimation_b200", "csrc"), "-o", str(lib), str(src)], check=True)
    L = C.CDLL(str(lib))
    rng = np.random.default_rng(7)
    for t in range(4000):
        W, H = int(rng.integers(4, 150)), int(rng.integers(4, 150))
        m = int(rng.choice([4, 40]))
        p0 = (int(rng.integers(-m, W + m)), int(rng.integers(-m, H + m)))
        p1 = p0 if t % 40 == 0 else (int(rng.integers(-m, W + m)), int(rng.integers(-m, H + m)))
        ref = np.zeros((H, W), np.uint8)
        cv2.line(ref, p0, p1, 255, 3)
        got = np.zeros((H, W), np.uint8)
        L.h_line(got.ctypes.data, W, H, p0[0], p0[1], p1[0], p1[1])
        assert np.array_equal(ref, got), ("line", W, H, p0, p1)
    for t in range(1500):
        W, H = int(rng.integers(4, 60)), int(rng.integers(4, 60))
        c = (int(rng.integers(-7, W + 8)), int(rng.integers(-7, H + 8)))
        ref = np.zeros((H, W), np.uint8)
        cv2.circle(ref, c, 3, 255, thickness=3, lineType=8, shift=0)
        got = np.zeros((H, W), np.uint8)
        L.h_circle3(got.ctypes.data, W, H, c[0], c[1])
        assert np.array_equal(ref, got), ("circle", W, H, c)


@pytest.mark.gpu
def test_draw_humans_matches_cv2_images(cuda):
    rng = np.random.default_rng(11)
    dec = tp.PoseDecoder(max_batch=1, max_upsample=(184, 216))
    for (h, w, nh) in ((368, 432, 3), (368, 656, 6), (120, 97, 12), (37, 53, 4), (736, 1312, 25)):
        humans = random_humans(rng, nh)
        img = rng.integers(0, 255, (h, w, 3), dtype=np.uint8)
        want = reference_draw(img.copy(), humans)
        got = dec.draw_humans(img.copy(), humans)
        assert np.array_equal(got, want), (h, w, nh, int((got != want).any(axis=2).sum()))
        out = dec.draw_humans(img, humans, imgcopy=True)
        assert out is not img and np.array_equal(out, want) and not np.array_equal(img, want)
    dec.close()


@pytest.mark.gpu
def test_draw_batch_on_device_and_from_decoded_tables(cuda):
    import torch
    from tfpose_b200 import synth
    rng = np.random.default_rng(12)
    x = synth.synth_batch(4, 46, 54, config_id=2, first_frame=900)
    dec = tp.PoseDecoder(max_batch=4, max_upsample=(184, 216))
    res = dec.decode(torch.from_numpy(x).cuda(), (184, 216))
    imgs = rng.integers(0, 255, (4, 368, 432, 3), dtype=np.uint8)
    want = np.stack([reference_draw(imgs[i].copy(), res.humans(i)) for i in range(4)])
    assert sum(len(res.humans(i)) for i in range(4)) > 0
    d_imgs = torch.from_numpy(imgs).cuda()
    dec.draw_decoded(d_imgs)                                   # straight from the device tables
    torch.cuda.synchronize()
    assert np.array_equal(d_imgs.cpu().numpy(), want)
    d2 = dec.draw_humans(torch.from_numpy(imgs).cuda(), res.all_humans())      # Human objects, device images
    assert np.array_equal(d2.cpu().numpy(), want)
    got = dec.draw_humans(imgs.copy(), res.all_humans())                        # host batch
    assert np.array_equal(got, want)
    dec.close()
    post = tp.TfPosePostProcessor(target_size=(432, 368))
    hs = post.inference_from_output(x[0], upsample_size=4.0)
    assert np.array_equal(post.draw_humans(imgs[0].copy(), hs), want[0])
    post.close()
