"""Golden vectors for the matcher produced by the REFERENCE'S OWN Frame / KeyFrame / MapPoint / ORBmatcher /
ORBVocabulary code (compiled in place: oracle/ref_orbmatcher.cc -> oracle/_ref/libref_orbslam.so; OpenCV routines from
cv2).  Run once in the authoring container:

    python tests/golden/make_ref_match_golden.py

  ref_match.npz   per scene s and frame f in {0,1}: the Frame's undistorted keypoints (x, y, angle, octave),
                  descriptors, valid flags (1 = good MapPoint, 2 = bad MapPoint, 0 = none), image bounds, the
                  FeatureVector of Frame::ComputeBoW (vocabulary: a synthetic k=3, L=5 tree written to a temp file),
                  the grid CSR; then the outputs of SearchForInitialization, WindowSearch, SearchByBoW (kf_f, kf_kf)
                  and GetFeaturesInArea queries, each with its parameters.
"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(HERE))

SFI = [(100, 0.9, 1), (40, 0.8, 0)]
WIN = [(200, 0, 2**31 - 1, 0.8, 1), (100, 1, 3, 0.8, 0)]
BOW = [("kf_f", 0.75, 1), ("kf_f", 0.6, 0), ("kf_kf", 0.75, 1)]


def main():
    import refscene as RS
    from oracle import bow_oracle as B
    from oracle import oracle as O
    from oracle import ref_frontend as RF
    assert O.build_ref() and RF.available()
    out = {}
    V = B.synth_vocabulary(k=3, L=5, seed=4, stop_frac=0.02)
    td = tempfile.mkdtemp()
    vpath = os.path.join(td, "voc.txt")
    B.save_text(vpath, **V)
    scenes = [(7, RS.CAM_SHIPPED, 400), (8, RS.CAM_PINHOLE, 300)]
    out["nscenes"] = np.array(len(scenes))
    rng = np.random.default_rng(3)
    for s, (seed, cam, nf) in enumerate(scenes):
        w = RF.World(vpath, nf, 1.2, 6, 1, 20)
        ims = RS.images(O, seed)
        fr = [w.frame(ims[0], RS.K_of(cam), cam[4:9], reset_static=True), w.frame(ims[1], RS.K_of(cam), cam[4:9])]
        out[f"s{s}_bounds"] = np.array(fr[0].bounds, np.int32)
        for f, rf in enumerate(fr):
            valid = RS.valid_flags(rf.n, 10 * s + f)
            rf.set_valid(valid)
            _, fv = rf.compute_bow()
            cs, it = rf.grid()
            p = f"s{s}_f{f}_"
            out[p + "x"] = rf.xu; out[p + "y"] = rf.yu; out[p + "angle"] = rf.kps["angle"]
            out[p + "octave"] = rf.kps["octave"]; out[p + "desc"] = rf.desc; out[p + "valid"] = valid
            out[p + "fv_node"], out[p + "fv_start"], out[p + "fv_items"] = fv
            out[p + "cell_start"] = cs; out[p + "grid_items"] = it
            q = np.stack([rng.uniform(-30, 350, 60), rng.uniform(-30, 270, 60), rng.uniform(1, 90, 60)], 1).astype(np.float32)
            lv = np.array([[(-1, -1), (1, 1), (0, 3)][i % 3] for i in range(60)], np.int32)
            res = [rf.features_in_area(float(a), float(b), float(c), int(l0), int(l1)) for (a, b, c), (l0, l1) in zip(q, lv)]
            out[p + "area_q"] = q; out[p + "area_lv"] = lv
            out[p + "area_start"] = np.concatenate([[0], np.cumsum([len(r) for r in res])]).astype(np.int32)
            out[p + "area_items"] = np.concatenate(res).astype(np.int32) if res else np.zeros(0, np.int32)
        prev = np.stack([fr[0].xu, fr[0].yu], 1)
        for k, (win, ratio, ori) in enumerate(SFI):
            n, m, p2 = RF.search_for_initialization(fr[0], fr[1], prev, win, ratio, ori)
            out[f"s{s}_sfi{k}_n"] = np.array(n); out[f"s{s}_sfi{k}_m"] = m; out[f"s{s}_sfi{k}_prev"] = p2
        for k, (win, l0, l1, ratio, ori) in enumerate(WIN):
            n, m = RF.window_search(fr[0], fr[1], win, l0, l1, ratio, ori)
            out[f"s{s}_win{k}_n"] = np.array(n); out[f"s{s}_win{k}_m"] = m
        for k, (variant, ratio, ori) in enumerate(BOW):
            n, m = RF.search_by_bow(fr[0], fr[1], variant, ratio, ori)
            out[f"s{s}_bow{k}_n"] = np.array(n); out[f"s{s}_bow{k}_m"] = m
            print(s, variant, n)
    np.savez_compressed(os.path.join(HERE, "ref_match.npz"), **out)
    print(os.path.getsize(os.path.join(HERE, "ref_match.npz")), "bytes")


if __name__ == "__main__":
    main()
