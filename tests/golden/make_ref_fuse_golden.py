"""Golden vectors for ORBmatcher::Fuse (both overloads, src/ORBmatcher.cc:1016-1265), SearchBySim3 (:1267-1505) and
SearchByProjection(F1, F2, window) (:519-594) produced by the REFERENCE'S OWN KeyFrame / MapPoint / ORBmatcher code
(oracle/ref_orbmatcher.cc -> oracle/_ref/libref_orbslam.so; scenes: tests/refscene.py::fuse_cases).  Run once in the
authoring container:

    python tests/golden/make_ref_fuse_golden.py

  ref_fuse.npz   per scene s and case c: the frames (undistorted keypoints, descriptors, bounds, MapPoint flags), the
                 projected points, the parameters and everything the reference's call left observable.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(HERE))

SCENES = [(41, "CAM_SHIPPED", 300), (93, "CAM_PINHOLE", 300)]


def put_frame(out, k, t):
    out[k + "x"], out[k + "y"], out[k + "angle"], out[k + "octave"] = t.x, t.y, t.angle, t.octave
    out[k + "desc"], out[k + "valid"] = t.desc, t.valid
    out[k + "bounds"] = np.array([t.min_x, t.min_y, t.max_x, t.max_y], np.float32)


def put_points(out, k, p):
    out[k + "u"], out[k + "v"], out[k + "radius"] = p.u, p.v, p.radius
    out[k + "min_level"], out[k + "max_level"], out[k + "desc"], out[k + "valid"] = p.min_level, p.max_level, p.desc, p.valid


def main():
    import refscene as RS
    from oracle import oracle as O
    from oracle import ref_frontend as RF
    assert O.build_ref() and RF.available()
    out = {"nscenes": np.array(len(SCENES))}
    for s, (seed, cam, nf) in enumerate(SCENES):
        nc = 0
        for c in RS.fuse_cases(O, RF, seed, getattr(RS, cam), nf):
            k = f"s{s}_c{nc}_"
            out[k + "kind"], out[k + "name"] = np.array(c["kind"]), np.array(c["name"])
            out[k + "n"] = np.array(c["ref_n"])
            if c["kind"] in ("fuse", "fuse_sim3"):
                put_frame(out, k + "t_", c["target"]); put_points(out, k + "p_", c["points"])
                out[k + "tvalid"], out[k + "th_dist"], out[k + "added"] = c["tvalid"], np.array(c["th_dist"]), c["added"]
                if c["kind"] == "fuse":
                    out[k + "went_bad"], out[k + "repl"] = c["went_bad"], c["repl"]
                else:
                    out[k + "target_went_bad"] = c["target_went_bad"]
            elif c["kind"] == "sim3":
                put_frame(out, k + "f1_", c["f1"]); put_frame(out, k + "f2_", c["f2"])
                put_points(out, k + "p12_", c["p12"]); put_points(out, k + "p21_", c["p21"])
                out[k + "pre"], out[k + "th_dist"], out[k + "m"] = c["pre"], np.array(c["th_dist"]), c["ref_m"]
            else:
                put_frame(out, k + "f1_", c["f1"]); put_frame(out, k + "f2_", c["f2"])
                out[k + "uv"], out[k + "m"] = c["uv"], c["ref_m"]
                out[k + "window"], out[k + "ratio"] = np.array(c["window"]), np.array(c["ratio"], np.float32)
            print(s, c["name"], c["ref_n"])
            nc += 1
        out[f"s{s}_ncases"] = np.array(nc)
    path = os.path.join(HERE, "ref_fuse.npz")
    np.savez_compressed(path, **out)
    print(os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
