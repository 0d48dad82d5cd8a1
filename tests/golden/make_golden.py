"""Generates the committed golden vectors (run once in the build container, where cv2 4.13 is importable):

  cv2_primitives.npz  outputs of the REAL OpenCV primitives the reference calls (cv::resize, cv::GaussianBlur,
                      cv::FAST + NMS, cv::fastAtan2) on a seeded image -> pins the oracle's primitive models on
                      machines without cv2 (the GPU box).
  extract_small.npz   oracle extractor output (cv2-pinned primitives + restated glue) on synth(120,160,seed) frames.
  match_small.npz     oracle matcher outputs on a small frame pair.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(HERE))


def main():
    import cv2
    from oracle import match_oracle as MO
    from oracle import oracle as O
    import matchdata as MD

    rng = np.random.default_rng(2026)
    img = O.synth(96, 128, 42)
    det20 = cv2.FastFeatureDetector_create(20, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    det7 = cv2.FastFeatureDetector_create(7, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    k20 = np.array([(k.pt[0], k.pt[1], k.response) for k in det20.detect(img)], np.int32)
    k7 = np.array([(k.pt[0], k.pt[1], k.response) for k in det7.detect(img)], np.int32)
    yx = rng.integers(-500, 500, (64, 2)).astype(np.float32)
    at = np.array([cv2.fastAtan2(float(y), float(x)) for y, x in yx], np.float32)
    np.savez_compressed(os.path.join(HERE, "cv2_primitives.npz"), img=img,
                        resize=cv2.resize(img, (107, 80), interpolation=cv2.INTER_LINEAR),
                        blur=cv2.GaussianBlur(img, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101),
                        fast20=k20, fast7=k7, atan_yx=yx, atan=at, cv2_version=np.array(cv2.__version__))

    ex = {}
    for seed, (nf, nl, th, bm, rm) in enumerate([(200, 4, 20, 0, 0), (300, 5, 12, 1, 1)]):
        im = O.synth(120, 160, 900 + seed)
        k, d = O.OracleExtractor(nf, 1.2, nl, 1, th, bm, rm)(im)
        ex[f"cfg{seed}"] = np.array([nf, nl, th, bm, rm, 900 + seed], np.int32)
        ex[f"kps{seed}"] = k
        ex[f"desc{seed}"] = d
    np.savez_compressed(os.path.join(HERE, "extract_small.npz"), **ex)

    f1, f2, shift = MD.make_pair(O, 7, 320, 240, 400)
    v1, v2 = MD.random_valid(f1.n, 0.7, 1), MD.random_valid(f2.n, 0.3, 2)
    prev = np.stack([f1.x, f1.y], 1).astype(np.float32)
    n_init, m_init = MO.search_for_initialization(f1, f2, prev, 100, 0.9, True)
    n_win, m_win = MO.window_search(MD.with_valid(f1, v1), f2, 60, 0.8, True)
    n_bow, m_bow = MO.search_by_bow(MD.with_valid(f1, v1), f2, "kf_f", 0.75, True)
    n_kk, m_kk = MO.search_by_bow(MD.with_valid(f1, v1), MD.with_valid(f2, 1 - v2), "kf_kf", 0.75, True)
    top2 = MO.knn2(f1.desc[:64], f2.desc)
    arr = lambda f, p: {p + k: getattr(f, k) for k in ("x", "y", "angle", "octave", "desc")}
    np.savez_compressed(os.path.join(HERE, "match_small.npz"), **arr(f1, "a_"), **arr(f2, "b_"), v1=v1, v2=v2,
                        a_fv0=f1.featvec[0], a_fv1=f1.featvec[1], a_fv2=f1.featvec[2], b_fv0=f2.featvec[0],
                        b_fv1=f2.featvec[1], b_fv2=f2.featvec[2], wh=np.array([320, 240], np.int32),
                        n=np.array([n_init, n_win, n_bow, n_kk], np.int32), m_init=m_init, prev=prev, m_win=m_win,
                        m_bow=m_bow, m_kk=m_kk, top2=top2)
    print("golden vectors written to", HERE)


if __name__ == "__main__":
    main()
