"""Generates the BoW golden vectors with the REFERENCE ITSELF (run once in the authoring container, where
oracle/_ref/libref_dbow2.so = the unmodified DBoW2 compiled from /root/reference exists):

  bow_voc_ref.txt   a vocabulary the reference's TemplatedVocabulary::create built (k-means++ on synthetic
                    descriptors, k=6, L=3, TF_IDF / L1_NORM like ORBvoc) and its saveToTextFile wrote; the trailing
                    newline is stripped (the reference's loader turns it into a node with an uninitialised parent).
  bow_golden.npz    descriptors + the reference's transform(features, BowVector, FeatureVector, levelsup) outputs for
                    levelsup in {4 (Frame::ComputeBoW, nid_level <= 0), 2, 1, 0}, L1 scores between the frames, and
                    FORB::distance on 512 random pairs.

    python tests/golden/make_bow_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)


def clustered(rng, n, cent, p):
    idx = rng.integers(0, len(cent), n)
    bits = np.unpackbits(cent[idx], axis=1) ^ (rng.random((n, 256)) < p)
    return np.packbits(bits, axis=1)


def main():
    from oracle import bow_oracle as B
    from oracle import oracle as O
    assert O.build_ref(), "needs /root/reference (authoring container)"
    rng = np.random.default_rng(77)
    cent = rng.integers(0, 256, (3000, 32), dtype=np.uint8)   # (few centroids -> empty k-means clusters -> the
    # reference's create() dereferences a released Mat; with 3000 it goes through)
    train = clustered(rng, 12 * 400, cent, 0.18).reshape(12, 400, 32)
    voc = B.RefVocabulary.create(train, 6, 3, weighting=0, scoring=0)
    path = os.path.join(HERE, "bow_voc_ref.txt")
    voc.save_text(path)
    txt = open(path).read().rstrip("\n")
    open(path, "w").write(txt)
    voc = B.RefVocabulary.load_text(path)   # what everybody else sees: the 6-significant-digit weights of the file
    out = dict(info=np.array([voc.k, voc.L, voc.scoring, voc.weighting, voc.nwords], np.int32))
    frames = [clustered(rng, n, cent, 0.09) for n in (500, 383, 1, 64)]
    frames[3][::3] = rng.integers(0, 256, frames[3][::3].shape, dtype=np.uint8)
    bows = []
    for f, d in enumerate(frames):
        out[f"desc{f}"] = d
        for lu in (4, 2, 1, 0):
            (bw, bv), (fn, fs, fi) = voc.transform(d, lu)
            if lu == 4:
                bows.append((bw, bv))
                out[f"bow_word{f}"] = bw; out[f"bow_weight{f}"] = bv
            out[f"fv_node{f}_{lu}"] = fn; out[f"fv_start{f}_{lu}"] = fs; out[f"fv_items{f}_{lu}"] = fi
    out["score"] = np.array([[voc.score(a, b) for b in bows] for a in bows], np.float64)
    a = rng.integers(0, 256, (512, 32), dtype=np.uint8); b = rng.integers(0, 256, (512, 32), dtype=np.uint8)
    b[:64] = a[:64]; b[:64, 5] ^= 0x11
    out["dist_a"] = a; out["dist_b"] = b; out["dist"] = B.ref_forb_distance(a, b)
    np.savez_compressed(os.path.join(HERE, "bow_golden.npz"), **out)
    print("vocabulary:", out["info"], "words;", os.path.getsize(path), "bytes of text;",
          os.path.getsize(os.path.join(HERE, "bow_golden.npz")), "bytes of vectors")


if __name__ == "__main__":
    main()
