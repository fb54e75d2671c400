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


def axis_table(ssize, dsize):
    """Per output index: first-tap source offset (tap k reads clamp(ofs-1+k)) and the four 11-bit coefficients.
    ``scale = 1/(dsize/ssize)`` in double as cv::resize -> hal::resize computes it."""
    scale = 1.0 / (float(dsize) / float(ssize))
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


def resize_cubic_u8(img, dst_w, dst_h):
    """``cv2.resize(img, (dst_w, dst_h), interpolation=cv2.INTER_CUBIC)`` for uint8 [H,W] or [H,W,C] (IPP off)."""
    squeeze = img.ndim == 2
    if squeeze:
        img = img[:, :, None]
    sh, sw, cn = img.shape
    xo, xa = axis_table(sw, dst_w)
    yo, ya = axis_table(sh, dst_h)
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


def get_scaled_img(npimg, target_size):
    """``_get_scaled_img(npimg, None)`` (estimator.py:417-424): resize to ``target_size = (w, h)`` unless equal."""
    if npimg.shape[:2] != (target_size[1], target_size[0]):
        npimg = resize_cubic_u8(npimg, target_size[0], target_size[1])
    return [npimg], [(0.0, 0.0, 1.0, 1.0)]
