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


PROJ_PATH = os.path.join(os.path.dirname(__file__), "golden", "ref_proj.npz")


def proj_cases():
    """tests/golden/ref_proj.npz (make_ref_proj_golden.py): the reference's SearchByProjection family (local map, last
    frame, relocalisation, Sim3) -> dicts with target MatchFrame, ProjectedPoints, parameters and reference outputs."""
    from orb_slam_b200.matcher import MatchFrame, ProjectedPoints
    g = np.load(PROJ_PATH)
    for s in range(int(g["nscenes"])):
        b = g[f"s{s}_bounds"]
        for c in range(int(g[f"s{s}_ncases"])):
            k = f"s{s}_c{c}_"
            t = MatchFrame(g[f"s{s}_x"], g[f"s{s}_y"], g[f"s{s}_angle"], g[f"s{s}_octave"], g[f"s{s}_desc"],
                           g[k + "tvalid"], float(b[0]), float(b[1]), float(b[2]), float(b[3]), None)
            ang = g[k + "angle"]
            p = ProjectedPoints(g[k + "u"], g[k + "v"], g[k + "radius"], g[k + "min_level"], g[k + "max_level"],
                                g[k + "desc"], ang if len(ang) else None, g[k + "valid"])
            variant, th_dist, ori = (int(v) for v in g[k + "params"])
            yield dict(name=f"s{s}:{str(g[k + 'name'])}", target=t, points=p, variant="local_map" if variant == 0 else "best",
                       th_dist=th_dist, ratio=float(g[k + "ratio"]), ori=bool(ori), ref_n=int(g[k + "n"]), ref_m=g[k + "m"])
