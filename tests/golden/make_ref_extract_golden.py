"""Golden vectors produced by the REFERENCE'S OWN extractor (run once in the authoring container): the unmodified
/root/reference/src/ORBextractor.cc compiled in place (oracle/_ref/libref_orbextractor.so) with its OpenCV library
calls served by the installed cv2 (see oracle/ref_extractor.py).  retainBest in its OpenCV >= 3.4 form (mode 1, what
cv2 4.13 itself does) and in its OpenCV 2.4 form (mode 0, the reference's era).

  ref_extract.npz:  for each case c: img{c}, params{c} = (nfeatures, nlevels, scoreType, fastTh, scaleFactor*1000),
                    kps{c}_{mode} (28-byte cv::KeyPoint records), desc{c}_{mode}

    python tests/golden/make_ref_extract_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

CASES = [  # (height, width, seed, nfeatures, scaleFactor, nlevels, scoreType, fastTh)
    (120, 160, 1, 60, 1.2, 4, 1, 20),
    (96, 128, 2, 30, 1.2, 3, 0, 20),       # Harris score
    (151, 233, 3, 150, 1.2, 8, 1, 20),     # odd size, all 8 levels
    (240, 320, 5, 500, 1.2, 6, 1, 20),     # quotas not reached everywhere
    (120, 160, 4, 80, 1.5, 4, 1, 12),
]


def main():
    import cv2
    from oracle import oracle as O
    from oracle import ref_extractor as RE
    assert O.build_ref() and RE.available(), "needs /root/reference and cv2 (authoring container)"
    out = {"cv2_version": np.array(cv2.__version__), "ncases": np.array(len(CASES))}
    for c, (h, w, seed, nf, sf, nl, st, th) in enumerate(CASES):
        img = O.synth(h, w, seed)
        if c == 4:   # low contrast: most cells take the threshold-7 retry (:609-614)
            img = (img.astype(np.int32) // 6 + 100).astype(np.uint8)
        out[f"img{c}"] = img
        out[f"params{c}"] = np.array([nf, nl, st, th, int(round(sf * 1000))], np.int32)
        R = RE.RefExtractor(nf, sf, nl, st, th)
        for mode in (1, 0):
            RE.set_retain_mode(mode)
            k, d = R(img)
            out[f"kps{c}_{mode}"] = k
            out[f"desc{c}_{mode}"] = d
            print(c, mode, len(k))
    RE.set_retain_mode(1)
    np.savez_compressed(os.path.join(HERE, "ref_extract.npz"), **out)
    print(os.path.getsize(os.path.join(HERE, "ref_extract.npz")), "bytes")


if __name__ == "__main__":
    main()
