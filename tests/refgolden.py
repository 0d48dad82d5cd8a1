"""Reader of tests/golden/ref_match.npz (vectors produced by the reference's own Frame/ORBmatcher code)."""
import os

import numpy as np

PATH = os.path.join(os.path.dirname(__file__), "golden", "ref_match.npz")
SFI = [(100, 0.9, 1), (40, 0.8, 0)]
WIN = [(200, 0, 2**31 - 1, 0.8, 1), (100, 1, 3, 0.8, 0)]
BOW = [("kf_f", 0.75, 1), ("kf_f", 0.6, 0), ("kf_kf", 0.75, 1)]


def load():
    return np.load(PATH)


def frame(g, s, f):
    from orb_slam_b200.matcher import MatchFrame
    p = f"s{s}_f{f}_"
    b = g[f"s{s}_bounds"]
    return MatchFrame(g[p + "x"], g[p + "y"], g[p + "angle"], g[p + "octave"], g[p + "desc"],
                      (g[p + "valid"] == 1).astype(np.uint8), float(b[0]), float(b[2]), float(b[1]), float(b[3]),
                      (g[p + "fv_node"], g[p + "fv_start"], g[p + "fv_items"]))
