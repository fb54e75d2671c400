"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy) of the image pre-processing in front of the network:

* ``quantize_img``   : ``TfPoseEstimator._quantize_img`` (tf_pose/estimator.py:383-388)
* ``resize_cubic_u8``: the ``cv2.resize(npimg, target_size, interpolation=cv2.INTER_CUBIC)`` of
  ``TfPoseEstimator._get_scaled_img(npimg, None)`` (tf_pose/estimator.py:417-424), the call ``inference`` makes
  when ``resize_to_default`` (estimator.py:534-535).

The resize lives in a third-party dependency of the reference (``opencv-python``, unpinned in requirements.txt;
4.13.0 in this image).  Restated here is OpenCV's own algorithm for 8-bit bicubic resizing (imgproc resize.cpp:
``interpolateCubic`` with A = -0.75 in float32, coefficients rounded to 11-bit fixed point, integer horizontal pass,
vertical pass in float32 on the 128-bit SIMD lanes with the scalar fixed-point tail for the last (width*cn) % 8
elements).  PINNED: bit-exact against ``cv2.resize`` itself with ``cv2.ipp.setUseIPP(False)`` over random sizes
(tests/test_preproc.py) and against tests/golden/preproc_cubic.npz.  With Intel IPP enabled (the wheel's default)
cv2 routes some sizes to closed-source ippicv code whose results differ by 1 LSB on ~4 % of the pixels; that
variant is not reproducible from any published algorithm and is not the bar.

Only tests/ may import this module; the product path (tf-pose-estimation_b200/) never does.
"""
import numpy as np

f32 = np.float32


def quantize_img(npimg):
    """estimator.py:383-388: ``(npimg + 1.0) / (2.0 / 2**8)`` then ``astype(np.uint8)`` (truncation)."""
    q = npimg + 1.0
    q /= (2.0 / 2 ** 8)
    return q.astype(np.uint8)


def _cubic_coeffs(x):
    """OpenCV ``interpolateCubic`` (float32, A = -0.75), evaluated operation by operation (no FMA)."""
    A = f32(-0.75)
    one = f32(1)
    x = f32(x)
    x1 = f32(x + one)
    c0 = f32(f32(f32(f32(f32(f32(A * x1) - f32(-3.75)) * x1) + f32(-6.0)) * x1) - f32(-3.0))
    c1 = f32(f32(f32(f32(f32(f32(1.25) * x) - f32(2.25)) * x) * x) + one)
    xm = f32(one - x)
    c2 = f32(f32(f32(f32(f32(f32(1.25) * xm) - f32(2.25)) * xm) * xm) + one)
    c3 = f32(f32(f32(one - c0) - c1) - c2)
    return np.array([c0, c1, c2, c3], dtype=f32)


def axis_table(ssize, dsize, f=None):
    """Per output index: first-tap source offset (tap k reads clamp(ofs-1+k)) and the four 11-bit coefficients.
    ``scale = 1/(dsize/ssize)`` in double as cv::resize -> hal::resize computes it; ``1/f`` when the size came
    from a scale factor (``dsize=None, fx=f``)."""
    scale = 1.0 / (float(dsize) / float(ssize)) if f is None else 1.0 / float(f)
    ofs = np.zeros(dsize, np.int64)
    coef = np.zeros((dsize, 4), np.int64)
    for d in range(dsize):
        fx = f32((d + 0.5) * scale - 0.5)
        s = int(np.floor(fx))
        fx = f32(fx - f32(s))
        c = _cubic_coeffs(fx)
        coef[d] = np.clip(np.rint(c * f32(2048)), -32768, 32767).astype(np.int64)   # saturate_cast<short>(cvRound)
        ofs[d] = s
    return ofs, coef


def resize_cubic_u8_fxfy(img, fx, fy):
    """``cv2.resize(img, dsize=None, fx=fx, fy=fy, interpolation=cv2.INTER_CUBIC)`` (estimator.py:428,452,482):
    dsize = round-half-even(src * f), sampling step 1/f."""
    return resize_cubic_u8(img, int(round(img.shape[1] * fx)), int(round(img.shape[0] * fy)), fx, fy)


def resize_cubic_u8(img, dst_w, dst_h, fx=None, fy=None):
    """``cv2.resize(img, (dst_w, dst_h), interpolation=cv2.INTER_CUBIC)`` for uint8 [H,W] or [H,W,C] (IPP off)."""
    squeeze = img.ndim == 2
    if squeeze:
        img = img[:, :, None]
    sh, sw, cn = img.shape
    xo, xa = axis_table(sw, dst_w, fx)
    yo, ya = axis_table(sh, dst_h, fy)
    S = img.astype(np.int64)
    H = np.zeros((sh, dst_w, cn), np.int64)
    for k in range(4):
        H += S[:, np.clip(xo - 1 + k, 0, sw - 1), :] * xa[:, k][None, :, None]
    rows = [np.clip(yo - 1 + k, 0, sh - 1) for k in range(4)]
    # vertical, SIMD lanes: float32, b_k = beta_k * 2^-22, acc = S3*b3; acc = S_k*b_k + acc (k = 2,1,0), separately
    # rounded multiply and add, round-half-even to int, saturating packs
    b = ya.astype(f32) * f32(1.0 / (2048.0 * 2048.0))
    acc = (H[rows[3]].astype(f32) * b[:, 3][:, None, None]).astype(f32)
    for k in (2, 1, 0):
        acc = ((H[rows[k]].astype(f32) * b[:, k][:, None, None]).astype(f32) + acc).astype(f32)
    out = np.clip(np.rint(acc), 0, 255).astype(np.uint8)
    # scalar tail of every row: the last (dst_w*cn) % 8 elements take the fixed-point cast
    tail = (dst_w * cn) % 8
    if tail:
        V = sum(H[rows[k]] * ya[:, k][:, None, None] for k in range(4))
        fixed = np.clip((V + (1 << 21)) >> 22, 0, 255).astype(np.uint8)
        flat = out.reshape(dst_h, dst_w * cn)
        flat[:, dst_w * cn - tail:] = fixed.reshape(dst_h, dst_w * cn)[:, dst_w * cn - tail:]
        out = flat.reshape(dst_h, dst_w, cn)
    return out[:, :, 0] if squeeze else out


def sliding_windows(width, height, win_w, win_h, overlap):
    """Window enumeration of the ``slidingwindow`` package as the reference vendors it
    (tf_pose/slidingwindow/SlidingWindow.py:100-146 ``generateForSize``, called through ``sw.generate`` at
    estimator.py:439,467).  PINNED: tests/golden/scaled_img.json holds window lists produced by that vendored code
    (make_golden.py scaled).  Returns [(x, y, w, h)] in the package's order (x outer, y inner)."""
    ww, wh = min(win_w, width), min(win_h, height)
    step_x, step_y = ww - int(np.floor(ww * overlap)), wh - int(np.floor(wh * overlap))
    last_x, last_y = width - ww, height - wh
    xs, ys = list(range(0, last_x + 1, step_x)), list(range(0, last_y + 1, step_y))
    if not xs or xs[-1] != last_x:
        xs.append(last_x)
    if not ys or ys[-1] != last_y:
        ys.append(last_y)
    return [(x, y, ww, wh) for x in xs for y in ys]


def crop_roi(npimg, target_size, ratio_x, ratio_y):
    """estimator.py:502-516 as written: the padded branch builds ``npblank`` and falls off the end (returns None)."""
    target_w, target_h = target_size
    h, w = npimg.shape[:2]
    x = max(int(w * ratio_x - .5), 0)
    y = max(int(h * ratio_y - .5), 0)
    cropped = npimg[y:y + target_h, x:x + target_w]
    if cropped.shape[1] < target_w or cropped.shape[0] < target_h:
        return None
    return cropped


def get_scaled_img(npimg, target_size, scale=None):
    """``_get_scaled_img(npimg, scale)`` (estimator.py:417-500), every form of ``scale``."""
    tw, th = target_size
    img_h, img_w = npimg.shape[:2]
    if scale is None:
        if npimg.shape[:2] != (th, tw):
            npimg = resize_cubic_u8(npimg, tw, th)
        return [npimg], [(0.0, 0.0, 1.0, 1.0)]

    def base_scale(s):
        return max(tw / float(img_h), th / float(img_w)) * s          # :418 as written (w over h, h over w)

    if isinstance(scale, float) or (isinstance(scale, tuple) and len(scale) == 2):
        s, step = (scale, 0.2) if isinstance(scale, float) else scale
        f = base_scale(s)
        npimg = resize_cubic_u8_fxfy(npimg, f, f)
        o_h, o_w = npimg.shape[:2]
        if o_h < th or o_w < tw:
            big = np.zeros((max(th, o_h), max(tw, o_w), 3), np.uint8)
            big[:o_h, :o_w, :] = npimg
            npimg = big
        rois, ratios = [], []
        for (x, y, w, h) in sliding_windows(npimg.shape[1], npimg.shape[0], tw, th, step):
            rois.append(npimg[y:y + h, x:x + w])
            ratios.append((float(x) / o_w, float(y) / o_h, float(w) / o_w, float(h) / o_h))
        return rois, ratios
    if isinstance(scale, tuple) and len(scale) == 3:
        f = base_scale(scale[2])
        npimg = resize_cubic_u8_fxfy(npimg, f, f)
        ratio_w, ratio_h = tw / float(npimg.shape[1]), th / float(npimg.shape[0])
        ratio_x, ratio_y = max(scale[0] - ratio_w / 2., 0.0), max(scale[1] - ratio_h / 2., 0.0)
        if ratio_x + ratio_w > 1.0:
            ratio_x = 1. - ratio_w
        if ratio_y + ratio_h > 1.0:
            ratio_y = 1. - ratio_h
        return [crop_roi(npimg, target_size, ratio_x, ratio_y)], [(ratio_x, ratio_y, ratio_w, ratio_h)]
    return None
