"""ctypes binding of ``libtfpose_decode.so`` (include/tfpose_decode.h).

The library is the product: if it is missing this module raises -- there is no
Python/NumPy/torch implementation of any decode stage to fall back to.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# TFPOSE_DECODE_LIB: another build of the same library (A/B timing of kernel variants); default = the in-tree build
LIB_PATH = os.environ.get("TFPOSE_DECODE_LIB") or os.path.join(_HERE, "libtfpose_decode.so")

NUM_PARTS = 18
OK, ERR_INVALID, ERR_CUDA, ERR_CAPACITY, ERR_STATE = 0, 1, 2, 3, 4
FLAG_SCORE_TIE, FLAG_FOUND_GT2 = 1, 2
FLAG_PEAK_OVERFLOW, FLAG_CAND_OVERFLOW, FLAG_ROW_OVERFLOW, FLAG_HUMAN_OVERFLOW = 4, 8, 16, 32
MODE_ASBUILT, MODE_UPSTREAM = 0, 1


class Config(C.Structure):
    """struct tfpd_config"""
    _fields_ = [
        ("device", C.c_int), ("max_batch", C.c_int), ("max_h", C.c_int), ("max_w", C.c_int),
        ("thresh_heat", C.c_float), ("thresh_vector_score", C.c_float), ("thresh_vector_cnt1", C.c_int),
        ("thresh_part_cnt", C.c_int), ("thresh_human_score", C.c_float), ("mode", C.c_int),
        ("blur_ksize", C.c_int), ("blur_sigma", C.c_float),
        ("max_peaks", C.c_int), ("max_cands", C.c_int), ("max_rows", C.c_int), ("max_humans", C.c_int),
    ]


class HumanRec(C.Structure):
    """struct tfpd_human"""
    _fields_ = [
        ("score", C.c_float), ("cid", C.c_int16 * NUM_PARTS), ("x", C.c_int16 * NUM_PARTS),
        ("y", C.c_int16 * NUM_PARTS), ("part_score", C.c_float * NUM_PARTS),
    ]


class FrameInfo(C.Structure):
    """struct tfpd_frame_info"""
    _fields_ = [("n_humans", C.c_int), ("n_peaks", C.c_int), ("n_cands", C.c_int), ("flags", C.c_int),
                ("near_ties", C.c_int), ("n_rows", C.c_int)]


class Dense(C.Structure):
    """struct tfpd_dense (device pointers)"""
    _fields_ = [("heat_up", C.c_void_p), ("paf_up", C.c_void_p), ("peaks", C.c_void_p)]


# numpy views of the two result structs
import numpy as _np  # noqa: E402

HUMAN_DTYPE = _np.dtype([("score", "<f4"), ("cid", "<i2", (NUM_PARTS,)), ("x", "<i2", (NUM_PARTS,)),
                         ("y", "<i2", (NUM_PARTS,)), ("part_score", "<f4", (NUM_PARTS,))], align=True)
INFO_DTYPE = _np.dtype([("n_humans", "<i4"), ("n_peaks", "<i4"), ("n_cands", "<i4"), ("flags", "<i4"),
                        ("near_ties", "<i4"), ("n_rows", "<i4")])
assert HUMAN_DTYPE.itemsize == C.sizeof(HumanRec) and INFO_DTYPE.itemsize == C.sizeof(FrameInfo)

# name -> (restype, argtypes); every symbol include/tfpose_decode.h declares
_P = C.c_void_p
_I = C.c_int
SIGNATURES = {
    "tfpd_default_config": (None, [C.POINTER(Config)]),
    "tfpd_create": (_I, [C.POINTER(Config), C.POINTER(_P)]),
    "tfpd_destroy": (None, [_P]),
    "tfpd_last_error": (C.c_char_p, [_P]),
    "tfpd_decode_device": (_I, [_P, _P, _I, _I, _I, _I, _I, _P, C.POINTER(Dense)]),
    "tfpd_decode_host": (_I, [_P, _P, _I, _I, _I, _I, _I, _P, _P, _P]),
    "tfpd_decode_host_async": (_I, [_P, _P, _I, _I, _I, _I, _I, _P, _P, _P]),
    "tfpd_fetch_humans": (_I, [_P, _P, _P]),
    "tfpd_fetch_peaks": (_I, [_P, _I, _I, _P, _P, _P, _P, C.POINTER(_I)]),
    "tfpd_fetch_connections": (_I, [_P, _I, _I, _P, _P, _P, _P, C.POINTER(_I)]),
    "tfpd_process_paf": (_I, [_P, _I, _I, _I, _P, _I, _I, _I, _P, _I, _I, _I, _P]),
    "tfpd_process_paf_device": (_I, [_P, _I, _I, _I, _P, _I, _I, _I, _P, _I, _I, _I, _P, _P]),
    "tfpd_peak_table_noalloc": (_I, [_P, _P, _I, _I, C.c_float, _I, _P, _P, _P, _P, _P, _P, C.c_char_p, C.c_size_t]),
    "tfpd_get_num_humans": (_I, [_P]),
    "tfpd_get_part_cid": (_I, [_P, _I, _I]),
    "tfpd_get_score": (C.c_float, [_P, _I]),
    "tfpd_get_part_x": (_I, [_P, _I]),
    "tfpd_get_part_y": (_I, [_P, _I]),
    "tfpd_get_part_score": (C.c_float, [_P, _I]),
    "tfpd_version": (_I, []),
    "tfpd_gauss_kernel": (_I, [_I, C.c_float, _P]),
    "tfpd_last_launch_count": (_I, [_P]),
    "tfpd_set_profiling": (_I, [_P, _I]),
    "tfpd_last_stage_ms": (_I, [_P, C.POINTER(C.c_float * 4)]),
    "tfpd_debug_k1_times": (_I, [_P, _P, _I]),
    "tfpd_set_option": (_I, [_P, C.c_char_p, _I]),
    "tfpd_draw_humans": (_I, [_P, _P, _I, _I, _I, _I, _P, _P, _I, _P]),
    "tfpd_draw_decoded": (_I, [_P, _P, _I, _I, _P]),
    "tfpd_resize_cubic_u8": (_I, [_P, _P, _I, _I, _I, _I, _I, _P, _I, _I, _P]),
    "tfpd_resize_cubic_u8_fxfy": (_I, [_P, _P, _I, _I, _I, _I, _I, C.c_double, C.c_double, _P, _I, _I, _P]),
    "tfpd_quantize_img": (_I, [_P, _P, _I, C.c_size_t, _P, _P]),
}

_lib = None


class DecodeLibraryMissing(ImportError):
    pass


def lib():
    """The loaded CDLL with prototypes set.  Raises if the CUDA library is not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise DecodeLibraryMissing(
                "%s not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(make -C tf-pose-estimation_b200/csrc). There is no CPU fallback." % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)          # AttributeError here == header/library mismatch
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


class DecodeError(RuntimeError):
    def __init__(self, code, message):
        RuntimeError.__init__(self, "tfpose_decode error %d: %s" % (code, message))
        self.code = code


class CapacityError(DecodeError):
    """A per-frame table overflowed; raise the matching ``max_*`` in the config.  ``result`` (when the
    error comes from ``PoseDecoder.fetch``) is the ``DecodeResult`` of the whole batch: frames whose
    ``info['flags']`` carry no overflow bit are complete."""
    result = None


def default_config(**overrides):
    cfg = Config()
    lib().tfpd_default_config(C.byref(cfg))
    for k, v in overrides.items():
        if not hasattr(cfg, k):
            raise TypeError("unknown tfpd_config field %r" % k)
        setattr(cfg, k, v)
    return cfg


def gauss_kernel(ksize=25, sigma=3.0):
    """float32 [ksize,ksize] smoother weights as the library computes them (host code)."""
    out = _np.empty((ksize, ksize), _np.float32)
    rc = lib().tfpd_gauss_kernel(ksize, float(sigma), out.ctypes.data)
    if rc != OK:
        raise DecodeError(rc, "tfpd_gauss_kernel(%d)" % ksize)
    return out
