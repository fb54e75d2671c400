"""ctypes front-end of the CPU oracle -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs may import this module.  It wraps

* ``liboracle_maps.so``  (oracle/maps_oracle.c : TF-op restatement, parity UNPINNED),
* ``liboracle_paf.so``   (oracle/paf_oracle.cpp: process_paf restatement, pinned
  against oracle/_ref and tests/golden),
* ``_ref/libpafprocess_{asbuilt,upstream}.so`` -- the reference's own
  tf_pose/pafprocess/pafprocess.cpp compiled by oracle/Makefile.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# TFPD_ORACLE_LIBDIR: another build of the same checker libraries (tools/sanitize_cpu.sh points it at the
# ASan/UBSan builds under oracle/_san); default = the in-tree build
_LIBDIR = os.environ.get("TFPD_ORACLE_LIBDIR") or _HERE
_F = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_D = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_I = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")


def build(force=False):
    """Compile the oracle (and, when /root/reference exists, oracle/_ref)."""
    need = force or not all(os.path.exists(os.path.join(_HERE, f))
                            for f in ("liboracle_maps.so", "liboracle_paf.so"))
    ref_missing = not os.path.exists(os.path.join(_HERE, "_ref", "libpafprocess_asbuilt.so"))
    if need or (ref_missing and os.path.isdir("/root/reference")):
        subprocess.run(["make", "-C", _HERE, "all"], check=True, stdout=subprocess.DEVNULL)


_maps = None
_paf = None


def maps_lib():
    global _maps
    if _maps is None:
        build()
        L = C.CDLL(os.path.join(_LIBDIR, "liboracle_maps.so"))
        L.oracle_gauss_kernel.argtypes = [C.c_int, C.c_double, _F, _D]
        L.oracle_split.argtypes = [_F, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _F]
        L.oracle_resize_area.argtypes = [_F, C.c_int, C.c_int, C.c_int, _F, C.c_int, C.c_int]
        L.oracle_blur_f32.argtypes = [_F, C.c_int, C.c_int, C.c_int, _F, C.c_int, _F]
        L.oracle_blur_f64.argtypes = [_F, C.c_int, C.c_int, C.c_int, _F, C.c_int, _D]
        L.oracle_nms.argtypes = [_F, C.c_int, C.c_int, C.c_int, _F, C.c_void_p]
        for f in ("oracle_gauss_kernel", "oracle_split", "oracle_resize_area", "oracle_blur_f32",
                  "oracle_blur_f64", "oracle_nms"):
            getattr(L, f).restype = None
        _maps = L
    return _maps


def paf_lib():
    global _paf
    if _paf is None:
        build()
        L = C.CDLL(os.path.join(_LIBDIR, "liboracle_paf.so"))
        L.paf_oracle_run.argtypes = [_F, _F, _F, C.c_int, C.c_int, C.c_int, C.c_int]
        L.paf_oracle_run.restype = C.c_void_p
        L.paf_oracle_run_cfg.argtypes = [_F, _F, _F, C.c_int, C.c_int, C.c_int, C.c_int,
                                         C.c_float, C.c_float, C.c_int, C.c_int, C.c_float]
        L.paf_oracle_run_cfg.restype = C.c_void_p
        L.paf_oracle_free.argtypes = [C.c_void_p]
        for f in ("paf_oracle_flags", "paf_oracle_num_peaks", "paf_oracle_num_humans",
                  "paf_oracle_num_rows_unfiltered"):
            getattr(L, f).argtypes = [C.c_void_p]
            getattr(L, f).restype = C.c_int
        for f in ("paf_oracle_num_conns", "paf_oracle_num_cands"):
            getattr(L, f).argtypes = [C.c_void_p, C.c_int]
            getattr(L, f).restype = C.c_int
        L.paf_oracle_get_peaks.argtypes = [C.c_void_p, _I, _I, _I, _F]
        L.paf_oracle_get_conns.argtypes = [C.c_void_p, C.c_int, _I, _I, _F]
        L.paf_oracle_get_cands.argtypes = [C.c_void_p, C.c_int, _I, _I, _F]
        L.paf_oracle_get_human_rows.argtypes = [C.c_void_p, _F]
        _paf = L
    return _paf


# --------------------------------------------------------------------------- maps

def gauss_kernel(kernlen=25, nsig=3.0):
    """(K32[kernlen,kernlen] float32, s1d[kernlen] float64) -- smoother.py:62-71."""
    k = np.empty((kernlen, kernlen), np.float32)
    s = np.empty(kernlen, np.float64)
    maps_lib().oracle_gauss_kernel(kernlen, float(nsig), k, s)
    return k, s


def split(x):
    """estimator.py:321-322: (heat[h,w,19], paf[h,w,38]) from X[h,w,57]."""
    x = np.ascontiguousarray(x, np.float32)
    h, w, c = x.shape
    heat = np.empty((h, w, 19), np.float32)
    paf = np.empty((h, w, c - 19), np.float32)
    maps_lib().oracle_split(x, h, w, c, 0, 19, heat)
    maps_lib().oracle_split(x, h, w, c, 19, c, paf)
    return heat, paf


def resize_area(img, oh, ow):
    img = np.ascontiguousarray(img, np.float32)
    ih, iw, c = img.shape
    out = np.empty((oh, ow, c), np.float32)
    maps_lib().oracle_resize_area(img, ih, iw, c, out, oh, ow)
    return out


def blur(up, k2d=None, f64=False):
    up = np.ascontiguousarray(up, np.float32)
    if k2d is None:
        k2d = gauss_kernel()[0]
    k2d = np.ascontiguousarray(k2d, np.float32)
    H, W, c = up.shape
    if f64:
        g = np.empty((H, W, c), np.float64)
        maps_lib().oracle_blur_f64(up, H, W, c, k2d, k2d.shape[0], g)
    else:
        g = np.empty((H, W, c), np.float32)
        maps_lib().oracle_blur_f32(up, H, W, c, k2d, k2d.shape[0], g)
    return g


def nms(g, want_margin=False):
    g = np.ascontiguousarray(g, np.float32)
    H, W, c = g.shape
    peaks = np.empty_like(g)
    if want_margin:
        margin = np.empty_like(g)
        maps_lib().oracle_nms(g, H, W, c, peaks, margin.ctypes.data)
        return peaks, margin
    maps_lib().oracle_nms(g, H, W, c, peaks, None)
    return peaks


def dense_maps(x, H, W, k2d=None):
    """The three dense tensors of one sess.run (estimator.py:536-539) for one frame
    ``x[h,w,57]``: (peaks[H,W,19], heat_up[H,W,19], paf_up[H,W,38], g[H,W,19])."""
    heat, paf = split(x)
    heat_up = resize_area(heat, H, W)
    paf_up = resize_area(paf, H, W)
    g = blur(heat_up, k2d)
    return nms(g), heat_up, paf_up, g


def fragile_pixels(g32, k_margin_rel=1e-5, thresh=0.05):
    """Near-tie report for a blurred map (SURVEY.md 8d parity gates): boolean map
    of pixels whose NMS decision or threshold decision could flip under a
    relative perturbation of ``k_margin_rel`` of g.  Only pixels that matter
    downstream (value near/above the heat threshold) are reported."""
    _, margin = nms(g32, want_margin=True)
    scale = np.maximum(np.abs(g32), 1e-30)
    near_nms = np.abs(margin) <= 2.0 * k_margin_rel * scale
    near_thr = np.abs(g32 - thresh) <= 2.0 * k_margin_rel * thresh
    relevant = g32 > thresh * (1.0 - 4.0 * k_margin_rel)
    # a pixel is only fragile for the peak list if it is (or nearly is) a local max
    is_or_nearly_max = margin >= -2.0 * k_margin_rel * scale
    return relevant[..., :18] & is_or_nearly_max[..., :18] & (near_nms[..., :18] | near_thr[..., :18])


# --------------------------------------------------------------------------- paf

class PafResult(object):
    """Everything process_paf decided for one frame."""
    __slots__ = ("peaks", "cands", "conns", "rows", "flags", "rows_unfiltered")

    def humans_as_tuples(self):
        """[(score, {part: (x, y, part_score)})] in human order."""
        out = []
        for row in self.rows:
            parts = {}
            for k in range(18):
                cid = int(row[k])
                if cid < 0:
                    continue
                parts[k] = (int(self.peaks["x"][cid]), int(self.peaks["y"][cid]), float(self.peaks["score"][cid]))
            out.append((float(np.float32(row[18]) / np.float32(row[19])), parts))
        return out


def process_paf(peaks, heat, paf, mode="asbuilt", sort="std", cfg=None):
    """The restatement (oracle/paf_oracle.cpp).  mode: 'asbuilt' | 'upstream';
    sort: 'std' (std::sort like the reference) | 'total' (score desc, idx1, idx2)."""
    peaks = np.ascontiguousarray(peaks, np.float32)
    heat = np.ascontiguousarray(heat, np.float32)
    paf = np.ascontiguousarray(paf, np.float32)
    H, W, _ = peaks.shape
    L = paf_lib()
    m = {"asbuilt": 0, "upstream": 1}[mode]
    s = {"std": 0, "total": 1}[sort]
    if cfg is None:
        h = L.paf_oracle_run(peaks, heat, paf, H, W, m, s)
    else:
        h = L.paf_oracle_run_cfg(peaks, heat, paf, H, W, m, s, *cfg)
    try:
        R = PafResult()
        n = L.paf_oracle_num_peaks(h)
        part, x, y = (np.empty(n, np.int32) for _ in range(3))
        score = np.empty(n, np.float32)
        if n:
            L.paf_oracle_get_peaks(h, part, x, y, score)
        R.peaks = {"part": part, "x": x, "y": y, "score": score}
        R.conns, R.cands = [], []
        for limb in range(19):
            for getter, counter, store in ((L.paf_oracle_get_conns, L.paf_oracle_num_conns, R.conns),
                                           (L.paf_oracle_get_cands, L.paf_oracle_num_cands, R.cands)):
                k = counter(h, limb)
                a, b = np.empty(k, np.int32), np.empty(k, np.int32)
                sc = np.empty(k, np.float32)
                if k:
                    getter(h, limb, a, b, sc)
                store.append((a, b, sc))
        nh = L.paf_oracle_num_humans(h)
        R.rows = np.empty((nh, 20), np.float32)
        if nh:
            L.paf_oracle_get_human_rows(h, R.rows)
        R.flags = L.paf_oracle_flags(h)
        R.rows_unfiltered = L.paf_oracle_num_rows_unfiltered(h)
        return R
    finally:
        L.paf_oracle_free(h)


class ReferencePafprocess(object):
    """The reference's own compiled pafprocess.cpp, exposed with the module-level
    API the reference's swig shim exposes (tf_pose/pafprocess/pafprocess.py:
    process_paf / get_num_humans / get_part_cid / get_score / get_part_x /
    get_part_y / get_part_score), bound through the C++ mangled names."""

    def __init__(self, variant="asbuilt"):
        path = os.path.join(_LIBDIR, "_ref", "libpafprocess_%s.so" % variant)
        if not os.path.exists(path):
            build()
        if not os.path.exists(path):
            raise FileNotFoundError(path + " (oracle/_ref is built from /root/reference by oracle/Makefile)")
        L = C.CDLL(path)
        self._process = getattr(L, "_Z11process_pafiiiPfiiiS_iiiS_")
        self._process.argtypes = [C.c_int, C.c_int, C.c_int, _F] * 3
        self._process.restype = C.c_int
        self.get_num_humans = getattr(L, "_Z14get_num_humansv")
        self.get_num_humans.restype = C.c_int
        self.get_part_cid = getattr(L, "_Z12get_part_cidii")
        self.get_part_cid.argtypes = [C.c_int, C.c_int]
        self.get_part_cid.restype = C.c_int
        self.get_score = getattr(L, "_Z9get_scorei")
        self.get_score.argtypes = [C.c_int]
        self.get_score.restype = C.c_float
        self.get_part_x = getattr(L, "_Z10get_part_xi")
        self.get_part_x.argtypes = [C.c_int]
        self.get_part_x.restype = C.c_int
        self.get_part_y = getattr(L, "_Z10get_part_yi")
        self.get_part_y.argtypes = [C.c_int]
        self.get_part_y.restype = C.c_int
        self.get_part_score = getattr(L, "_Z14get_part_scorei")
        self.get_part_score.argtypes = [C.c_int]
        self.get_part_score.restype = C.c_float
        self._keep = None

    def process_paf(self, peaks, heat_mat, paf_mat):
        # swig's IN_ARRAY3 typemap converts to contiguous float32 (pafprocess_wrap.cpp:4935-4968)
        p = np.ascontiguousarray(peaks, np.float32)
        h = np.ascontiguousarray(heat_mat, np.float32)
        f = np.ascontiguousarray(paf_mat, np.float32)
        for a in (p, h, f):
            if a.ndim != 3:
                raise ValueError("Array must have 3 dimensions.")
        self._keep = (p, h, f)
        return self._process(p.shape[0], p.shape[1], p.shape[2], p, h.shape[0], h.shape[1], h.shape[2], h,
                             f.shape[0], f.shape[1], f.shape[2], f)

    def humans_as_tuples(self):
        out = []
        for hid in range(self.get_num_humans()):
            parts = {}
            for k in range(18):
                cid = self.get_part_cid(hid, k)
                if cid < 0:
                    continue
                parts[k] = (self.get_part_x(cid), self.get_part_y(cid), float(self.get_part_score(cid)))
            out.append((float(self.get_score(hid)), parts))
        return out


def decode_frame(x, H, W, mode="asbuilt", sort="std", k2d=None):
    """Whole path for one frame on the CPU: dense maps + restated process_paf."""
    peaks, heat_up, paf_up, g = dense_maps(x, H, W, k2d)
    return process_paf(peaks, heat_up, paf_up, mode=mode, sort=sort), (peaks, heat_up, paf_up, g)
