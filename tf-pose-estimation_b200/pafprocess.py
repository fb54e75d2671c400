"""Drop-in for the reference's swig module ``tf_pose.pafprocess.pafprocess`` (seam 2).

Same module-level API as tf_pose/pafprocess/pafprocess.py:110-130 (prototypes
tf_pose/pafprocess/pafprocess.h:53-59): ``process_paf(peaks, heat_mat, paf_mat)`` then the
getters.  Like the reference the results of the last call live in module state
(pafprocess.cpp:12-13) -- here one lazily created GPU context per process.  The work is done
by the CUDA kernels behind ``tfpd_process_paf``; nothing is computed in Python.
"""
from . import _lib
from .decoder import PoseDecoder

_decoder = None
_settings = {"device": 0, "mode": "asbuilt", "max_peaks": 4096, "max_cands": 8192, "max_rows": 1024,
             "max_humans": 256}


def configure(**kwargs):
    """Change device / mode ('asbuilt' | 'upstream') / capacities; recreates the context."""
    global _decoder
    _settings.update(kwargs)
    if _decoder is not None:
        _decoder.close()
        _decoder = None


def _ctx():
    global _decoder
    if _decoder is None:
        s = dict(_settings)
        _decoder = PoseDecoder(max_batch=1, max_upsample=(8191, 16383), device=s.pop("device"), mode=s.pop("mode"), **s)
    return _decoder


def process_paf(peaks, heat_mat, paf_mat):
    return _ctx().process_paf(peaks, heat_mat, paf_mat)


def get_num_humans():
    d = _ctx()
    return int(d._L.tfpd_get_num_humans(d._ctx))


def get_part_cid(human_id, part_id):
    d = _ctx()
    return int(d._L.tfpd_get_part_cid(d._ctx, int(human_id), int(part_id)))


def get_score(human_id):
    d = _ctx()
    return float(d._L.tfpd_get_score(d._ctx, int(human_id)))


def get_part_x(cid):
    d = _ctx()
    return int(d._L.tfpd_get_part_x(d._ctx, int(cid)))


def get_part_y(cid):
    d = _ctx()
    return int(d._L.tfpd_get_part_y(d._ctx, int(cid)))


def get_part_score(cid):
    d = _ctx()
    return float(d._L.tfpd_get_part_score(d._ctx, int(cid)))


# constants the swig module exports (pafprocess.h:6-15)
THRESH_HEAT = 0.05
THRESH_VECTOR_SCORE = 0.05
THRESH_VECTOR_CNT1 = 6
THRESH_PART_CNT = 4
THRESH_HUMAN_SCORE = 0.3
NUM_PART = 18
STEP_PAF = 10
COCOPAIRS_SIZE = 19
FLAG_SCORE_TIE = _lib.FLAG_SCORE_TIE
