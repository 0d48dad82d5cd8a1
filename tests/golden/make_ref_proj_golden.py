"""Golden vectors for the SearchByProjection family (local map, last frame, relocalisation, Sim3) produced by
the REFERENCE'S OWN Frame / KeyFrame / MapPoint / ORBmatcher code (oracle/ref_orbmatcher.cc ->
oracle/_ref/libref_orbslam.so; scenes: tests/refscene.py::projection_cases).  Run once in the authoring container:

    python tests/golden/make_ref_proj_golden.py

  ref_proj.npz   per scene s: the target frame (undistorted keypoints, descriptors, bounds); per case c of the scene:
                 the target's "holds a MapPoint" flags, the projected points (u, v, radius, level window, descriptor,
                 angle, valid), the parameters (variant, th_dist, ratio, ori) and the reference's (n, matches).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(HERE))

SCENES = [(31, "CAM_SHIPPED", 300), (72, "CAM_PINHOLE", 300)]


def main():
    import refscene as RS
    from oracle import oracle as O
    from oracle import ref_frontend as RF
    assert O.build_ref() and RF.available()
    out = {"nscenes": np.array(len(SCENES))}
    for s, (seed, cam, nf) in enumerate(SCENES):
        ncases = 0
        for c in RS.projection_cases(O, RF, seed, getattr(RS, cam), nf):
            t, p = c["target"], c["points"]
            if ncases == 0:
                out[f"s{s}_x"], out[f"s{s}_y"], out[f"s{s}_angle"] = t.x, t.y, t.angle
                out[f"s{s}_octave"], out[f"s{s}_desc"] = t.octave, t.desc
                out[f"s{s}_bounds"] = np.array([t.min_x, t.min_y, t.max_x, t.max_y], np.float32)
            k = f"s{s}_c{ncases}_"
            out[k + "name"] = np.array(c["name"])
            out[k + "tvalid"] = t.valid
            out[k + "u"], out[k + "v"], out[k + "radius"] = p.u, p.v, p.radius
            out[k + "min_level"], out[k + "max_level"], out[k + "desc"] = p.min_level, p.max_level, p.desc
            out[k + "angle"] = p.angle if p.angle is not None else np.zeros(0, np.float32)
            out[k + "valid"] = p.valid
            out[k + "params"] = np.array([0 if c["variant"] == "local_map" else 1, c["th_dist"], int(c["ori"])], np.int32)
            out[k + "ratio"] = np.array(c["ratio"], np.float32)
            out[k + "n"] = np.array(c["ref_n"]); out[k + "m"] = c["ref_m"]
            print(s, c["name"], c["ref_n"])
            ncases += 1
        out[f"s{s}_ncases"] = np.array(ncases)
    np.savez_compressed(os.path.join(HERE, "ref_proj.npz"), **out)
    print(os.path.getsize(os.path.join(HERE, "ref_proj.npz")), "bytes")


if __name__ == "__main__":
    main()
