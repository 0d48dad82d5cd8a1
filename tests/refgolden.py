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


FUSE_PATH = os.path.join(os.path.dirname(__file__), "golden", "ref_fuse.npz")


def fuse_cases():
    """tests/golden/ref_fuse.npz (make_ref_fuse_golden.py): the reference's Fuse (both overloads), SearchBySim3 and
    SearchByProjection(F1, F2, window) -> the dicts of tests/refscene.py::fuse_cases."""
    from orb_slam_b200.matcher import MatchFrame, ProjectedPoints
    g = np.load(FUSE_PATH)

    def fr(k):
        b = g[k + "bounds"]
        return MatchFrame(g[k + "x"], g[k + "y"], g[k + "angle"], g[k + "octave"], g[k + "desc"], g[k + "valid"],
                          float(b[0]), float(b[1]), float(b[2]), float(b[3]), None)

    def pts(k):
        return ProjectedPoints(g[k + "u"], g[k + "v"], g[k + "radius"], g[k + "min_level"], g[k + "max_level"],
                               g[k + "desc"], None, g[k + "valid"])

    for s in range(int(g["nscenes"])):
        for c in range(int(g[f"s{s}_ncases"])):
            k = f"s{s}_c{c}_"
            kind = str(g[k + "kind"])
            d = dict(kind=kind, name=f"s{s}:{str(g[k + 'name'])}", ref_n=int(g[k + "n"]))
            if kind in ("fuse", "fuse_sim3"):
                d.update(target=fr(k + "t_"), points=pts(k + "p_"), tvalid=g[k + "tvalid"], th_dist=int(g[k + "th_dist"]),
                         added=g[k + "added"])
                if kind == "fuse":
                    d.update(went_bad=g[k + "went_bad"], repl=g[k + "repl"])
                else:
                    d.update(target_went_bad=g[k + "target_went_bad"])
            elif kind == "sim3":
                d.update(f1=fr(k + "f1_"), f2=fr(k + "f2_"), p12=pts(k + "p12_"), p21=pts(k + "p21_"), pre=g[k + "pre"],
                         th_dist=int(g[k + "th_dist"]), ref_m=g[k + "m"])
            else:
                d.update(f1=fr(k + "f1_"), f2=fr(k + "f2_"), uv=g[k + "uv"], window=int(g[k + "window"]),
                         ratio=float(g[k + "ratio"]), ref_m=g[k + "m"])
            yield d
