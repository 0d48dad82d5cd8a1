"""N2 oracle pinning: oracle/bow_oracle.cc against (a) golden vectors produced by the reference's own DBoW2
(tests/golden/make_bow_golden.py) and (b), where oracle/_ref/libref_dbow2.so exists, the compiled reference live over
every scoring x weighting combination (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1126-1259)."""
import os

import numpy as np
import pytest

from oracle import bow_oracle as B
from oracle import match_oracle as MO

GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLD, "bow_golden.npz"))


@pytest.fixture(scope="module")
def voc():
    return B.Vocabulary.load_text(os.path.join(GOLD, "bow_voc_ref.txt"))


def same_f64(a, b):
    return np.array_equal(np.asarray(a, np.float64).view(np.int64), np.asarray(b, np.float64).view(np.int64))


def test_vocabulary_header(gold, voc):
    assert [voc.k, voc.L, voc.scoring, voc.weighting, voc.nwords] == list(gold["info"])
    parent, desc, weight, leaf = voc.arrays()
    assert voc.nnodes == 259 and parent[0] == -1 and int(leaf.sum()) == voc.nwords
    assert np.all(parent[1:] < np.arange(1, voc.nnodes))


@pytest.mark.parametrize("frame", [0, 1, 2, 3])
def test_transform_matches_reference_golden(gold, voc, frame):
    d = gold[f"desc{frame}"]
    for lu in (4, 2, 1, 0):
        (bw, bv), (fn, fs, fi) = voc.transform(d, lu)
        assert np.array_equal(bw, gold[f"bow_word{frame}"])
        assert same_f64(bv, gold[f"bow_weight{frame}"])          # bit-exact doubles
        assert np.array_equal(fn, gold[f"fv_node{frame}_{lu}"])
        assert np.array_equal(fs, gold[f"fv_start{frame}_{lu}"])
        assert np.array_equal(fi, gold[f"fv_items{frame}_{lu}"])


def test_l1_score_matches_reference_golden(gold, voc):
    bows = [voc.transform(gold[f"desc{f}"], 4)[0] for f in range(4)]
    s = np.array([[B.score_l1(a, b) for b in bows] for a in bows])
    assert same_f64(s, gold["score"])
    assert np.allclose(np.diag(s), 1.0, atol=1e-12)


def test_forb_distance_golden(gold):
    # FORB::distance (FORB.cpp:81-101) == ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1794-1810)
    d = [MO.descriptor_distance(a, b) for a, b in zip(gold["dist_a"], gold["dist_b"])]
    assert np.array_equal(d, gold["dist"])


def test_text_round_trip_and_blank_line_ends_file(tmp_path):
    V = B.synth_vocabulary(k=4, L=3, seed=5, stop_frac=0.1)
    p = str(tmp_path / "v.txt")
    B.save_text(p, **V, trailing_newline=True)   # the reference's own files end with a newline
    O = B.Vocabulary.load_text(p)
    parent, desc, weight, leaf = O.arrays()
    assert np.array_equal(parent, V["parent"]) and np.array_equal(desc[1:], V["desc"][1:])
    assert same_f64(weight, V["weight"]) and np.array_equal(leaf, V["is_leaf"])
    O2 = B.Vocabulary.from_arrays(V["k"], V["L"], 0, 0, parent, desc, weight, leaf)
    d = np.random.default_rng(0).integers(0, 256, (200, 32), dtype=np.uint8)
    a, b = O.transform(d, 2), O2.transform(d, 2)
    assert all(np.array_equal(x, y) for x, y in zip(a[0] + a[1], b[0] + b[1]))


def test_bad_vocabularies_rejected(tmp_path):
    V = B.synth_vocabulary(k=3, L=2, seed=1)
    bad = V["parent"].copy(); bad[2] = 7                       # child before its parent
    with pytest.raises(ValueError):
        B.Vocabulary.from_arrays(3, 2, 0, 0, bad, V["desc"], V["weight"], V["is_leaf"])
    leaf = V["is_leaf"].copy(); leaf[1] = 1                    # "leaf" that has children
    with pytest.raises(ValueError):
        B.Vocabulary.from_arrays(3, 2, 0, 0, V["parent"], V["desc"], V["weight"], leaf)
    p = tmp_path / "bad.txt"; p.write_text("99 3 0 0\n")
    with pytest.raises(ValueError):
        B.Vocabulary.load_text(str(p))


def test_stopped_words_and_empty_input():
    V = B.synth_vocabulary(k=4, L=2, seed=9, stop_frac=0.5)
    O = B.Vocabulary.from_arrays(**{k: V[k] for k in ("k", "L", "scoring", "weighting", "parent", "desc", "weight",
                                                      "is_leaf")})
    (bw, bv), (fn, fs, fi) = O.transform(np.zeros((0, 32), np.uint8), 1)
    assert len(bw) == 0 and len(fn) == 0 and list(fs) == [0]
    d = V["desc"][V["is_leaf"] == 1]
    word, w, node = O.descend(d, 1)
    assert np.array_equal(word, np.arange(len(d)))             # every leaf descriptor descends to its own word
    (bw, bv), (fn, fs, fi) = O.transform(d, 1)
    assert np.array_equal(bw, word[w > 0]) and np.array_equal(np.sort(fi), np.flatnonzero(w > 0))
    assert abs(bv.sum() - 1.0) < 1e-12


@pytest.mark.skipif(not B.have_ref(), reason="oracle/_ref/libref_dbow2.so not built (needs /root/reference)")
@pytest.mark.parametrize("scoring", range(6))
@pytest.mark.parametrize("weighting", range(4))
def test_oracle_equals_compiled_reference(tmp_path, scoring, weighting):
    rng = np.random.default_rng(100 + scoring * 4 + weighting)
    for ragged in (False, True):
        V = B.synth_vocabulary(k=5, L=3, seed=scoring * 10 + weighting, scoring=scoring, weighting=weighting,
                               ragged=ragged, stop_frac=0.1)
        p = str(tmp_path / f"v{int(ragged)}.txt")
        B.save_text(p, **V)
        R = B.RefVocabulary.load_text(p)
        O = B.Vocabulary.load_text(p)
        assert (R.k, R.L, R.scoring, R.weighting, R.nwords) == (O.k, O.L, O.scoring, O.weighting, O.nwords)
        leaves = V["desc"][V["is_leaf"] == 1]
        for n in (0, 1, 7, 300):
            d = rng.integers(0, 256, (n, 32), dtype=np.uint8)
            if n > 1:
                d[::2] = leaves[rng.integers(0, len(leaves), len(d[::2]))]
            for lu in (0, 1, 2, 4):
                (rw, rv), rf = R.transform(d, lu)
                (ow, ov), of = O.transform(d, lu)
                assert np.array_equal(rw, ow) and same_f64(rv, ov)
                if not ragged or V["L"] - lu <= 1:   # ragged + deep nid_level: the reference's *nid is uninitialised
                    assert all(np.array_equal(x, y) for x, y in zip(rf, of))


@pytest.mark.skipif(not B.have_ref(), reason="oracle/_ref/libref_dbow2.so not built (needs /root/reference)")
def test_descriptor_distance_equals_compiled_forb():
    rng = np.random.default_rng(4)
    a = rng.integers(0, 256, (2048, 32), dtype=np.uint8); b = rng.integers(0, 256, (2048, 32), dtype=np.uint8)
    assert np.array_equal(B.ref_forb_distance(a, b), [MO.descriptor_distance(x, y) for x, y in zip(a, b)])
