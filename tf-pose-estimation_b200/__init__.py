"""B200-native heatmap/PAF decode for tf-pose-estimation network outputs.

Drop-in for the post-processing of ``TfPoseEstimator.inference``
(reference tf_pose/estimator.py:321-337,518-553 + tf_pose/pafprocess): Python
host code over a ctypes C-ABI (include/tfpose_decode.h) over hand-written
sm_100a CUDA kernels.  There is no CPU fallback: every decode call needs the
built ``libtfpose_decode.so`` and a CUDA device.
"""
from .common import CocoPart, CocoPairs, CocoPairsRender, CocoPairsNetwork, CocoColors  # noqa: F401
from .humans import Human, BodyPart  # noqa: F401
from .decoder import PoseDecoder, DecodeResult, upsample_size_for, humans_from_table  # noqa: F401
from .estimator import PoseEstimator, TfPosePostProcessor  # noqa: F401
from . import writers  # noqa: F401
from ._lib import DecodeError, CapacityError, DecodeLibraryMissing  # noqa: F401

__all__ = ["CocoPart", "CocoPairs", "CocoPairsRender", "CocoPairsNetwork", "CocoColors", "Human", "BodyPart",
           "PoseDecoder", "DecodeResult", "PoseEstimator", "TfPosePostProcessor", "upsample_size_for",
           "humans_from_table", "DecodeError", "CapacityError", "DecodeLibraryMissing"]
