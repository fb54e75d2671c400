"""Part / limb tables of the decode path (reference: tf_pose/common.py:13-32,98-110
and tf_pose/pafprocess/pafprocess.h:16-24).  Data only -- the reference module's
image helpers and TF constants are outside the decode path."""
from enum import Enum


class CocoPart(Enum):
    # tf_pose/common.py:13-32 ; index == heat-map channel
    Nose = 0
    Neck = 1
    RShoulder = 2
    RElbow = 3
    RWrist = 4
    LShoulder = 5
    LElbow = 6
    LWrist = 7
    RHip = 8
    RKnee = 9
    RAnkle = 10
    LHip = 11
    LKnee = 12
    LAnkle = 13
    REye = 14
    LEye = 15
    REar = 16
    LEar = 17
    Background = 18


NUM_PARTS = 18          # pafprocess.h:11
NUM_LIMBS = 19          # pafprocess.h:15
NUM_HEAT_CHANNELS = 19  # estimator.py:321
NUM_PAF_CHANNELS = 38   # estimator.py:322
NUM_NET_CHANNELS = NUM_HEAT_CHANNELS + NUM_PAF_CHANNELS

# limb -> (partA, partB): common.py:98-101 == pafprocess.h:21-24
CocoPairs = [
    (1, 2), (1, 5), (2, 3), (3, 4), (5, 6), (6, 7), (1, 8), (8, 9), (9, 10), (1, 11),
    (11, 12), (12, 13), (1, 0), (0, 14), (14, 16), (0, 15), (15, 17), (2, 16), (5, 17),
]
CocoPairsRender = CocoPairs[:-2]   # common.py:102 (ear-shoulder limbs are not drawn)

# limb -> (PAF x-channel, PAF y-channel): pafprocess.h:16-19 (common.py:103-106 has it commented out)
CocoPairsNetwork = [
    (12, 13), (20, 21), (14, 15), (16, 17), (22, 23), (24, 25), (0, 1), (2, 3), (4, 5),
    (6, 7), (8, 9), (10, 11), (28, 29), (30, 31), (34, 35), (32, 33), (36, 37), (18, 19), (26, 27),
]

# common.py:108-110
CocoColors = [[255, 0, 0], [255, 85, 0], [255, 170, 0], [255, 255, 0], [170, 255, 0], [85, 255, 0],
              [0, 255, 0], [0, 255, 85], [0, 255, 170], [0, 255, 255], [0, 170, 255], [0, 85, 255],
              [0, 0, 255], [85, 0, 255], [170, 0, 255], [255, 0, 255], [255, 0, 170], [255, 0, 85]]
