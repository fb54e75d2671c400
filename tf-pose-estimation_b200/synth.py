"""Seeded synthetic network outputs ``X[N,h,w,57]`` for the decode path.

There are no weights or datasets on the box, so bench and tests decode maps
drawn from skeletons the way the reference's ground-truth generator draws them
(tf_pose/pose_dataset.py:107-214): one Gaussian per joint on the 18 part
channels (sigma = 1 cell == the GT generator's 8 px at stride 8,
pose_dataset.py:127-149), background channel ``clip(1 - max, 0, 1)`` (:120) and
a unit vector inside a +-1-cell band along every limb on the PAF channel pair
``COCOPAIRS_NET[limb]`` (:181-214, channel table tf_pose/pafprocess/pafprocess.h:16-19).
Joint positions are sub-cell and N(0, 0.005) noise is added to all 57 channels so
exact ties have measure zero.
"""
import numpy as np

from .common import CocoPairs, CocoPairsNetwork

# 18-joint template, neck at the origin, x right / y down, unit = body size.
_TEMPLATE = np.array([
    (0.00, -0.10),   # Nose
    (0.00, 0.00),    # Neck
    (-0.10, 0.00),   # RShoulder
    (-0.13, 0.15),   # RElbow
    (-0.14, 0.30),   # RWrist
    (0.10, 0.00),    # LShoulder
    (0.13, 0.15),    # LElbow
    (0.14, 0.30),    # LWrist
    (-0.06, 0.30),   # RHip
    (-0.065, 0.50),  # RKnee
    (-0.07, 0.70),   # RAnkle
    (0.06, 0.30),    # LHip
    (0.065, 0.50),   # LKnee
    (0.07, 0.70),    # LAnkle
    (-0.02, -0.12),  # REye
    (0.02, -0.12),   # LEye
    (-0.04, -0.11),  # REar
    (0.04, -0.11),   # LEar
], dtype=np.float64)


def frame_seed(config_id, frame):
    """Seed rule of SURVEY.md 8d: s = 1000*config + frame."""
    return 1000 * int(config_id) + int(frame)


def synth_frame(h, w, seed, people=None, noise=0.005):
    """One frame ``[h,w,57]`` float32.  ``people=None`` draws P ~ U{1..6}."""
    rng = np.random.default_rng(seed)
    n_people = int(rng.integers(1, 7)) if people is None else int(people)
    ys, xs = np.mgrid[0:h, 0:w].astype(np.float64)
    heat = np.zeros((h, w, 18), np.float64)
    paf_sum = np.zeros((h, w, 38), np.float64)
    paf_cnt = np.zeros((h, w, 19), np.float64)
    base = 0.9 * h
    for _ in range(n_people):
        scale = rng.uniform(0.5, 1.0) * base
        tmpl = _TEMPLATE * scale
        lo, hi = tmpl.min(0), tmpl.max(0)
        # neck placed uniformly so that the whole body stays inside the map
        nx = rng.uniform(1.0 - lo[0], max(1.0 - lo[0] + 1e-3, w - 2.0 - hi[0]))
        ny = rng.uniform(1.0 - lo[1], max(1.0 - lo[1] + 1e-3, h - 2.0 - hi[1]))
        joints = tmpl + np.array([nx, ny]) + rng.normal(0.0, 0.2, size=tmpl.shape)
        joints[:, 0] = np.clip(joints[:, 0], 0.0, w - 1.0)
        joints[:, 1] = np.clip(joints[:, 1], 0.0, h - 1.0)
        for k in range(18):
            d2 = (xs - joints[k, 0]) ** 2 + (ys - joints[k, 1]) ** 2
            np.maximum(heat[:, :, k], np.exp(-d2 / 2.0), out=heat[:, :, k])
        for limb, (pa, pb) in enumerate(CocoPairs):
            ax, ay = joints[pa]
            bx, by = joints[pb]
            vx, vy = bx - ax, by - ay
            length = float(np.hypot(vx, vy))
            if length < 1e-6:
                continue
            ux, uy = vx / length, vy / length
            along = (xs - ax) * ux + (ys - ay) * uy
            perp = np.abs((xs - ax) * uy - (ys - ay) * ux)
            inside = (along >= -1.0) & (along <= length + 1.0) & (perp <= 1.0)
            chx, chy = CocoPairsNetwork[limb]
            paf_sum[:, :, chx][inside] += ux
            paf_sum[:, :, chy][inside] += uy
            paf_cnt[:, :, limb][inside] += 1.0
    out = np.zeros((h, w, 57), np.float64)
    out[:, :, :18] = heat
    out[:, :, 18] = np.clip(1.0 - heat.max(axis=2), 0.0, 1.0)
    for limb in range(19):
        chx, chy = CocoPairsNetwork[limb]
        cnt = np.maximum(paf_cnt[:, :, limb], 1.0)
        out[:, :, 19 + chx] = paf_sum[:, :, chx] / cnt
        out[:, :, 19 + chy] = paf_sum[:, :, chy] / cnt
    out += rng.normal(0.0, noise, size=out.shape)
    return out.astype(np.float32)


def synth_batch(n, h, w, config_id=2, people=None, first_frame=0, noise=0.005):
    """``[n,h,w,57]`` float32; frame i uses seed ``1000*config_id + first_frame + i``."""
    out = np.empty((n, h, w, 57), np.float32)
    for i in range(n):
        out[i] = synth_frame(h, w, frame_seed(config_id, first_frame + i), people=people, noise=noise)
    return out
