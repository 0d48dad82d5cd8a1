"""Scenes for the compiled-reference pins (test infrastructure): a pair of shifted views, built into the reference's own
Frame objects (oracle/ref_frontend.py) and, from what those hold, into the flat MatchFrame records the oracle and the
CUDA matcher take."""
from __future__ import annotations

import numpy as np

CAM_SHIPPED = (268.9633, 269.9858, 157.6087, 114.6369, -0.4157, 0.2624, 0.0, 0.0, -0.1178)   # Data/Settings.yaml
CAM_PINHOLE = (300.0, 300.0, 160.0, 120.0, 0.0, 0.0, 0.0, 0.0, 0.0)                           # k1 == 0: no undistortion


def images(oracle, seed, h=240, w=320, shift=(11, 6)):
    big = oracle.synth(h + 64, w + 64, seed)
    a = np.ascontiguousarray(big[32:32 + h, 32:32 + w])
    b = np.ascontiguousarray(big[32 + shift[1]:32 + shift[1] + h, 32 + shift[0]:32 + shift[0] + w])
    return a, b


def K_of(cam):
    return [cam[0], 0, cam[2], 0, cam[1], cam[3], 0, 0, 1]


def valid_flags(n, seed):
    rng = np.random.default_rng(seed)
    v = (rng.random(n) < 0.7).astype(np.uint8)
    v[rng.random(n) < 0.05] = 2          # a MapPoint whose isBad() is true
    return v


def match_frame(rf, valid, featvec):
    """MatchFrame from what the reference Frame holds (undistorted keypoints, bounds)."""
    from orb_slam_b200.matcher import MatchFrame
    b = rf.bounds
    return MatchFrame(rf.xu, rf.yu, rf.kps["angle"], rf.kps["octave"], rf.desc, (np.asarray(valid) == 1).astype(np.uint8),
                      float(b[0]), float(b[2]), float(b[1]), float(b[3]), featvec)
