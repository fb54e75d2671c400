"""Image pre-processing in front of the network, SURVEY.md 8f rank 4: the bicubic resize of
``TfPoseEstimator._get_scaled_img(npimg, None)`` (tf_pose/estimator.py:417-424) and ``_quantize_img`` (:383-388).

Checker = oracle/preproc_oracle.py, itself pinned (CPU tests) against tests/golden/preproc_cubic.npz (outputs of
``cv2.resize(..., INTER_CUBIC)`` with IPP off, made by tests/golden/make_golden.py preproc) and, where cv2 is
importable, against cv2 live on random sizes.  GPU tests compare the CUDA kernels through the C-ABI, bit for bit."""
import hashlib
import os

import numpy as np
import pytest

import tfpose_b200 as tp
from oracle import preproc_oracle as PO

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = np.load(os.path.join(HERE, "golden", "preproc_cubic.npz"))


def golden_image(seed, sh, sw, cn):
    return np.random.default_rng(9000 + seed).integers(0, 256, (sh, sw, cn), dtype=np.uint8)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_oracle_matches_golden_cv2_outputs():
    for (seed, sh, sw, cn, dw, dh), want in zip(GOLD["cases"].tolist(), GOLD["sha256"].tolist()):
        out = PO.resize_cubic_u8(golden_image(seed, sh, sw, cn), dw, dh)
        assert out.shape == (dh, dw, cn)
        got = sha(out[:, :, 0] if cn == 1 else out)     # cv2 drops a single channel axis
        assert got == want, (seed, sh, sw, cn, dw, dh)
    assert np.array_equal(PO.resize_cubic_u8(golden_image(7, 7, 9, 3), 20, 20), GOLD["small"])


def test_oracle_matches_cv2_live_on_random_sizes():
    cv2 = pytest.importorskip("cv2")
    had = cv2.ipp.useIPP()
    cv2.ipp.setUseIPP(False)      # OpenCV's own code; the closed-source IPP path differs by 1 LSB on some sizes
    try:
        rng = np.random.default_rng(77)
        for _ in range(25):
            sh, sw, dw, dh = (int(v) for v in rng.integers(3, 200, 4))
            cn = int(rng.choice([1, 3, 4]))
            img = rng.integers(0, 256, (sh, sw, cn), dtype=np.uint8)
            ref = cv2.resize(img, (dw, dh), interpolation=cv2.INTER_CUBIC).reshape(dh, dw, cn)
            assert np.array_equal(PO.resize_cubic_u8(img, dw, dh), ref), (sh, sw, cn, dw, dh)
    finally:
        cv2.ipp.setUseIPP(had)


def test_oracle_identity_and_quantize():
    img = golden_image(3, 20, 30, 3)
    assert PO.get_scaled_img(img, (30, 20))[0][0] is img               # same size: returned untouched (:420)
    assert PO.get_scaled_img(img, (15, 10))[0][0].shape == (10, 15, 3)
    assert PO.get_scaled_img(img, (15, 10))[1] == [(0.0, 0.0, 1.0, 1.0)]
    x = np.array([-1.0, -0.5, 0.0, 0.4999, 0.99], np.float32)
    assert PO.quantize_img(x).tolist() == [0, 64, 128, 191, 254]


def test_oracle_fxfy_matches_cv2_live():
    cv2 = pytest.importorskip("cv2")
    had = cv2.ipp.useIPP()
    cv2.ipp.setUseIPP(False)
    try:
        rng = np.random.default_rng(78)
        for i in range(16):
            sh, sw = (int(v) for v in rng.integers(5, 200, 2))
            fx = float(rng.uniform(0.3, 2.5))
            fy = fx if i % 2 == 0 else float(rng.uniform(0.3, 2.5))
            img = rng.integers(0, 256, (sh, sw, 3), dtype=np.uint8)
            ref = cv2.resize(img, dsize=None, fx=fx, fy=fy, interpolation=cv2.INTER_CUBIC)
            out = PO.resize_cubic_u8_fxfy(img, fx, fy)
            assert out.shape == ref.shape and np.array_equal(out, ref), (sh, sw, fx, fy)
    finally:
        cv2.ipp.setUseIPP(had)


def scaled_golden():
    import json
    with open(os.path.join(HERE, "golden", "scaled_img.json")) as fh:
        return json.load(fh)


def scaled_cases():
    return scaled_golden()["cases"]


def check_scaled(case, rois, ratios):
    assert len(rois) == len(case["rois"]) and len(ratios) == len(case["ratios"]), case["scale"]
    for r, want in zip(rois, case["rois"]):
        if want is None:
            assert r is None
        else:
            assert list(r.shape) == want[0] and sha(r) == want[1], case["scale"]
    for got, want in zip(ratios, case["ratios"]):
        assert [float(v) for v in got] == want, case["scale"]


def test_oracle_get_scaled_img_matches_the_reference_method():
    """tests/golden/scaled_img.json holds what the reference's own _get_scaled_img / _crop_roi returned (real cv2,
    IPP off) on top of the reference's own vendored slidingwindow package (tf_pose/slidingwindow)."""
    for case in scaled_cases():
        scale = tuple(case["scale"]) if isinstance(case["scale"], list) else case["scale"]
        img = golden_image(case["seed"], case["h"], case["w"], 3)
        check_scaled(case, *PO.get_scaled_img(img, tuple(case["target"]), scale))
    assert PO.sliding_windows(500, 184, 216, 184, 0.2) == [(0, 0, 216, 184), (173, 0, 216, 184), (284, 0, 216, 184)]
    assert PO.sliding_windows(100, 100, 216, 184, 0.2) == [(0, 0, 100, 100)]


def test_window_enumeration_matches_the_vendored_slidingwindow_package():
    """Window lists of the reference's vendored ``slidingwindow.generateForSize``
    (tf_pose/slidingwindow/SlidingWindow.py:100-146; golden made by make_golden.py scaled): the oracle's restatement
    and the product's ``_sliding_windows`` must enumerate the same (x, y, w, h) in the same order."""
    from tfpose_b200.estimator import _sliding_windows
    wins = scaled_golden()["sliding_windows"]
    assert len(wins) >= 12
    for c in wins:
        args = (c["width"], c["height"], c["win_w"], c["win_h"], c["overlap"])
        want = [tuple(w) for w in c["windows"]]
        assert PO.sliding_windows(*args) == want, args
        assert _sliding_windows(*args) == want, args


@pytest.fixture(scope="module")
def decoder(cuda):
    d = tp.PoseDecoder(max_batch=1, max_upsample=(184, 216))
    yield d
    d.close()


@pytest.mark.gpu
@pytest.mark.parametrize("path", [0, 1, 2, 3])       # automatic (staged strip, else strip) / tiled / one thread per output pixel / strip
def test_resize_cubic_matches_oracle_and_golden(decoder, path):
    decoder.set_option("resize_path", path)
    try:
        for (seed, sh, sw, cn, dw, dh), want in zip(GOLD["cases"].tolist(), GOLD["sha256"].tolist()):
            img = golden_image(seed, sh, sw, cn)
            out = decoder.resize_cubic(img, (dw, dh))
            assert out.shape == (dh, dw, cn) and out.dtype == np.uint8
            assert np.array_equal(out, PO.resize_cubic_u8(img, dw, dh)), (seed, sh, sw, cn, dw, dh)
            assert sha(out[:, :, 0] if cn == 1 else out) == want
        big = golden_image(11, 900, 700, 3)                         # x7 / x9 down-scale: tall source tiles
        assert np.array_equal(decoder.resize_cubic(big, (100, 100)), PO.resize_cubic_u8(big, 100, 100))
    finally:
        decoder.set_option("resize_path", 0)
    g = golden_image(5, 50, 61, 1)[:, :, 0]
    assert decoder.resize_cubic(g, (43, 37)).shape == (37, 43)         # 2-D in, 2-D out like cv2


@pytest.mark.gpu
def test_resize_cubic_staged_form_on_odd_shapes(decoder):
    """The automatic path takes the shared-memory staged strip kernel whenever the source rows are 4-byte aligned:
    row tails that take the fixed-point cast (dst_w * cn not a multiple of 8), up- and down-scaling in either
    direction, 1 / 2 / 3 / 4 channels, windows that run over every image border, tiles narrower than a CTA and
    strips shorter than 32 rows -- against the CPU oracle (pinned on cv2) and against the plain strip kernel."""
    import torch
    shapes = [(37, 64, 1, 91, 53), (120, 160, 3, 431, 97), (480, 640, 3, 433, 367), (97, 128, 2, 45, 200),
              (64, 48, 4, 47, 31), (300, 400, 3, 100, 77), (33, 20, 3, 5, 7), (5, 4, 3, 300, 40), (2, 4, 1, 1, 1),
              (61, 256, 3, 256, 61), (200, 1024, 3, 700, 33)]
    for i, (sh, sw, cn, dw, dh) in enumerate(shapes):
        assert (sw * cn) % 4 == 0
        img = golden_image(100 + i, sh, sw, cn)
        want = PO.resize_cubic_u8(img, dw, dh)
        decoder.set_option("resize_path", 0)
        got = decoder.resize_cubic(img, (dw, dh))
        decoder.set_option("resize_path", 3)
        plain = decoder.resize_cubic(img, (dw, dh))
        decoder.set_option("resize_path", 0)
        got = got.reshape(want.shape)
        assert np.array_equal(got, want), (sh, sw, cn, dw, dh, int(np.abs(got.astype(int) - want.astype(int)).max()))
        assert np.array_equal(plain.reshape(want.shape), want)
    # a batch on the device, extremes that overshoot on both sides
    rng = np.random.default_rng(9)
    imgs = rng.integers(0, 256, (5, 240, 320, 3), dtype=np.uint8)
    imgs[1, ::2] = 0
    imgs[1, 1::2] = 255
    imgs[2, :, ::2] = 255
    imgs[2, :, 1::2] = 0
    out = decoder.resize_cubic(torch.from_numpy(imgs).cuda(), (217, 163)).cpu().numpy()
    for i in range(len(imgs)):
        assert np.array_equal(out[i], PO.resize_cubic_u8(imgs[i], 217, 163)), i


@pytest.mark.gpu
def test_resize_cubic_batch_on_device(decoder):
    import torch
    rng = np.random.default_rng(5)
    imgs = rng.integers(0, 256, (6, 480, 640, 3), dtype=np.uint8)
    imgs[1] = 255
    imgs[2] = 0
    imgs[3, ::2] = 0
    imgs[3, 1::2] = 255                                                 # overshoot: saturation on both sides
    out = decoder.resize_cubic(torch.from_numpy(imgs).cuda(), (432, 368))
    torch.cuda.synchronize()
    out = out.cpu().numpy()
    for i in range(len(imgs)):
        assert np.array_equal(out[i], PO.resize_cubic_u8(imgs[i], 432, 368)), i
    assert np.array_equal(decoder.resize_cubic(imgs, (432, 368)), out)  # host-pointer call, same bytes
    for shift in (1, 2, 3):                                             # source not 4-byte aligned: word loads realign
        flat = torch.zeros(imgs[:2].size + 8, dtype=torch.uint8, device="cuda")
        view = flat[shift:shift + imgs[:2].size].view(2, 480, 640, 3)
        view.copy_(torch.from_numpy(imgs[:2]))
        assert np.array_equal(decoder.resize_cubic(view, (432, 368)).cpu().numpy(), out[:2]), shift


@pytest.mark.gpu
@pytest.mark.parametrize("target", [(432, 368), (656, 368), (1312, 736)])      # network inputs of BASELINE.json's configs
def test_resize_cubic_camera_frames_to_network_inputs(decoder, target):
    import torch
    rng = np.random.default_rng(target[0])
    # smooth content + noise + saturated patches, 1920x1080 like a camera frame
    yy, xx = np.mgrid[0:1080, 0:1920]
    base = 127.5 + 100.0 * np.sin(xx / 37.0) * np.cos(yy / 23.0)
    imgs = np.stack([np.clip(base[:, :, None] + rng.normal(0, 30, (1080, 1920, 3)), 0, 255).astype(np.uint8)
                     for _ in range(2)])
    imgs[0, 100:200, 300:500] = 255
    imgs[1, 900:, :64] = 0
    out = decoder.resize_cubic(torch.from_numpy(imgs).cuda(), target).cpu().numpy()
    for i in range(2):
        assert np.array_equal(out[i], PO.resize_cubic_u8(imgs[i], target[0], target[1])), (target, i)


@pytest.mark.gpu
def test_quantize_and_scaled_img_mirror(decoder, cuda):
    import torch
    rng = np.random.default_rng(6)
    x = rng.uniform(-1.0, 1.0, (368, 432, 3)).astype(np.float32)
    x[0, 0] = (-1.0, 0.0, 0.99999994)
    want = PO.quantize_img(x)
    assert np.array_equal(decoder.quantize_img(x), want)
    assert np.array_equal(decoder.quantize_img(torch.from_numpy(x).cuda()).cpu().numpy(), want)
    odd = x.ravel()[:1001]                                              # count not a multiple of 4: scalar tail
    assert np.array_equal(decoder.quantize_img(odd), PO.quantize_img(odd))
    shifted = torch.from_numpy(x).cuda().ravel()[1:1000].contiguous()
    assert np.array_equal(decoder.quantize_img(shifted.clone()).cpu().numpy(), PO.quantize_img(x.ravel()[1:1000]))
    assert np.array_equal(decoder.quantize_img(torch.from_numpy(x).cuda().ravel()[1:1000]).cpu().numpy(),
                          PO.quantize_img(x.ravel()[1:1000]))           # 4-byte aligned only: unvectorised path
    post = tp.TfPosePostProcessor(target_size=(432, 368))
    try:
        img = golden_image(0, 480, 640, 3)
        rois, ratios = post._get_scaled_img(img, None)
        assert ratios == [(0.0, 0.0, 1.0, 1.0)] and np.array_equal(rois[0], PO.resize_cubic_u8(img, 432, 368))
        same = golden_image(1, 368, 432, 3)
        assert post._get_scaled_img(same, None)[0][0] is same
        assert np.array_equal(post._quantize_img(x), want)
    finally:
        post.close()


@pytest.mark.gpu
def test_preproc_rejects_bad_arguments(decoder):
    with pytest.raises(ValueError):
        decoder.resize_cubic(np.zeros((4, 4, 3), np.float32), (8, 8))
    with pytest.raises(Exception):
        decoder.resize_cubic(np.zeros((4, 4, 5), np.uint8), (8, 8))    # 5 channels: TFPD_ERR_INVALID
    with pytest.raises(ValueError):
        decoder.quantize_img(np.zeros((4, 4), np.float64))


@pytest.mark.gpu
def test_inference_shaped_chain_stays_on_the_device(cuda, oracle):
    """The steps of ``TfPoseEstimator.inference`` + ``draw_humans`` around the (out-of-scope) network, all on the
    device and chained on one stream: camera frame -> bicubic resize to the network input (estimator.py:534-535) ->
    [network output: synthetic] -> decode (:536-551) -> skeletons drawn into the resized frame (:391-415).  Checked
    against the CPU chain: resize oracle (= cv2's bytes), decode oracle, cv2 drawing."""
    import torch
    from helpers import humans_equal
    from test_draw import reference_draw
    from tfpose_b200 import synth
    rng = np.random.default_rng(21)
    n = 3
    frames = rng.integers(0, 256, (n, 480, 640, 3), dtype=np.uint8)
    x = synth.synth_batch(n, 46, 54, config_id=2, first_frame=700)
    dec = tp.PoseDecoder(max_batch=n, max_upsample=(184, 216))
    try:
        d_net_in = dec.resize_cubic(torch.from_numpy(frames).cuda(), (432, 368))
        res = dec.decode(torch.from_numpy(x).cuda(), (184, 216))
        dec.draw_decoded(d_net_in)
        torch.cuda.synchronize()
        got = d_net_in.cpu().numpy()
        assert sum(len(res.humans(i)) for i in range(n)) > 0
        for i in range(n):
            R, _ = oracle.decode_frame(x[i], 184, 216)
            if int(res.info["near_ties"][i]) == 0:
                ok, why = humans_equal(res.humans(i), R.humans_as_tuples(), 184, 216)
                assert ok, (i, why)
            want = reference_draw(PO.resize_cubic_u8(frames[i], 432, 368), res.humans(i))
            assert np.array_equal(got[i], want), i
    finally:
        dec.close()


@pytest.mark.gpu
def test_get_scaled_img_every_form_matches_the_reference_method(cuda):
    for case in scaled_cases():
        scale = tuple(case["scale"]) if isinstance(case["scale"], list) else case["scale"]
        post = tp.TfPosePostProcessor(target_size=tuple(case["target"]))
        try:
            check_scaled(case, *post._get_scaled_img(golden_image(case["seed"], case["h"], case["w"], 3), scale))
            assert post._get_scaled_img(golden_image(1, 8, 8, 3), "bad") is None
        finally:
            post.close()


@pytest.mark.gpu
def test_resize_fxfy_rejects_a_size_cv2_would_not_make(decoder):
    import ctypes as C
    img = golden_image(2, 40, 50, 3)
    out = decoder.resize_cubic(img, None, fx=1.3, fy=0.7)
    assert np.array_equal(out, PO.resize_cubic_u8_fxfy(img, 1.3, 0.7))
    dst = np.empty((1, 29, 65, 3), np.uint8)                # cv2 makes 28 x 65 from 40 x 50 at fy = 0.7
    rc = decoder._L.tfpd_resize_cubic_u8_fxfy(decoder._ctx, img.ctypes.data, 0, 1, 40, 50, 3, C.c_double(1.3),
                                              C.c_double(0.7), dst.ctypes.data, 29, 65, None)
    assert rc == 1                                           # TFPD_ERR_INVALID
