"""N2 parity: the device BoW transform (orbm_vocabulary_* / orbm_frameset_compute_bow, through the C ABI) against
the golden vectors the reference's own DBoW2 produced and against the pinned CPU oracle.  Word / node ids and feature
lists are integers, weights and scores are doubles compared BIT for bit."""
import os

import numpy as np
import pytest

import matchdata as MD

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden")
KEYS = ("k", "L", "scoring", "weighting", "parent", "desc", "weight", "is_leaf")


def same_f64(a, b):
    return np.array_equal(np.asarray(a, np.float64).view(np.int64), np.asarray(b, np.float64).view(np.int64))


def same_transform(got, ref):
    (gw, gv), gf = got
    (rw, rv), rf = ref
    return np.array_equal(gw, rw) and same_f64(gv, rv) and all(np.array_equal(x, y) for x, y in zip(gf, rf))


@pytest.fixture(scope="module")
def B(oracle):
    from oracle import bow_oracle
    return bow_oracle


@pytest.fixture(scope="module")
def matcher():
    from orb_slam_b200 import Matcher
    return Matcher(0)


def test_golden_vocabulary_and_transforms(matcher):
    """ORBvoc.txt-format file written by the reference's saveToTextFile; outputs by the reference's transform."""
    from orb_slam_b200 import ORBVocabulary
    g = np.load(os.path.join(GOLD, "bow_golden.npz"))
    voc = ORBVocabulary.loadFromTextFile(matcher, os.path.join(GOLD, "bow_voc_ref.txt"))
    assert [voc.k, voc.L, voc.scoring, voc.weighting, voc.nwords] == list(g["info"]) and voc.nnodes == 259
    for f in range(4):
        for lu in (4, 2, 1, 0):
            (bw, bv), (fn, fs, fi) = voc.transform(g[f"desc{f}"], lu)
            assert np.array_equal(bw, g[f"bow_word{f}"]) and same_f64(bv, g[f"bow_weight{f}"]), (f, lu)
            assert np.array_equal(fn, g[f"fv_node{f}_{lu}"]) and np.array_equal(fs, g[f"fv_start{f}_{lu}"])
            assert np.array_equal(fi, g[f"fv_items{f}_{lu}"])


def test_golden_scores_on_resident_frames(matcher):
    from orb_slam_b200 import MatchFrame, ORBVocabulary
    g = np.load(os.path.join(GOLD, "bow_golden.npz"))
    voc = ORBVocabulary.loadFromTextFile(matcher, os.path.join(GOLD, "bow_voc_ref.txt"))
    frames = []
    for f in range(4):
        d = g[f"desc{f}"]
        z = np.zeros(len(d), np.float32)
        frames.append(MatchFrame(z, z, z, np.zeros(len(d), np.int32), d, max_x=640.0, max_y=480.0))
    fs = matcher.upload(frames)
    fs.compute_bow(voc, 4)
    for f in range(4):
        w, v = fs.bow_vector(f)
        assert np.array_equal(w, g[f"bow_word{f}"]) and same_f64(v, g[f"bow_weight{f}"])
        nid, st, it = fs.feature_vector(f)
        assert np.array_equal(nid, g[f"fv_node{f}_4"]) and np.array_equal(st, g[f"fv_start{f}_4"])
        assert np.array_equal(it, g[f"fv_items{f}_4"])
    i1, i2 = np.divmod(np.arange(16), 4)
    assert same_f64(voc.score(fs, fs, i1, i2).reshape(4, 4), g["score"])


@pytest.mark.parametrize("scoring", range(6))
@pytest.mark.parametrize("weighting", range(4))
def test_every_scoring_and_weighting(B, matcher, scoring, weighting):
    from orb_slam_b200 import ORBVocabulary
    rng = np.random.default_rng(scoring * 4 + weighting)
    for ragged in (False, True):
        V = B.synth_vocabulary(k=5, L=3, seed=scoring * 10 + weighting + 1, scoring=scoring, weighting=weighting,
                               ragged=ragged, stop_frac=0.1)
        O = B.Vocabulary.from_arrays(**{k: V[k] for k in KEYS})
        G = ORBVocabulary.from_arrays(matcher, **{k: V[k] for k in KEYS})
        assert (G.k, G.L, G.scoring, G.weighting, G.nwords, G.nnodes) == (O.k, O.L, O.scoring, O.weighting, O.nwords,
                                                                         O.nnodes)
        leaves = V["desc"][V["is_leaf"] == 1]
        for n in (0, 1, 7, 300):
            d = rng.integers(0, 256, (n, 32), dtype=np.uint8)
            if n > 1:
                d[::2] = leaves[rng.integers(0, len(leaves), len(d[::2]))]
            for lu in (0, 1, 2, 4):
                assert same_transform(G.transform(d, lu), O.transform(d, lu)), (ragged, n, lu)


def test_large_frame_sorts_in_global_scratch_and_k10(B, matcher):
    """> 4096 keys leaves the shared-memory sort; k = 10, L = 4 is the ORBvoc branching factor (odd tail child)."""
    from orb_slam_b200 import ORBVocabulary
    V = B.synth_vocabulary_fast(k=10, L=4, seed=3)
    V["weight"][-50:] = 0.0   # stopped words
    O = B.Vocabulary.from_arrays(**{k: V[k] for k in KEYS})
    G = ORBVocabulary.from_arrays(matcher, **{k: V[k] for k in KEYS})
    rng = np.random.default_rng(8)
    leaves = V["desc"][V["is_leaf"] == 1]
    for n in (4096, 4097, 9000):
        d = leaves[rng.integers(0, len(leaves), n)].copy()
        flip = rng.integers(0, 256, n)
        d[np.arange(n), flip >> 3] ^= (1 << (flip & 7)).astype(np.uint8)
        d[:: 5] = d[0]      # many features of one word: the repeated `+= w`
        assert same_transform(G.transform(d, 2), O.transform(d, 2)), n


def test_text_file_equals_arrays_and_errors(B, matcher, tmp_path):
    from orb_slam_b200 import OrbmError, ORBVocabulary
    V = B.synth_vocabulary(k=4, L=3, seed=12, stop_frac=0.05)
    p = str(tmp_path / "voc.txt")
    B.save_text(p, **V, trailing_newline=True)
    G1 = ORBVocabulary.loadFromTextFile(matcher, p)
    G2 = ORBVocabulary.from_arrays(matcher, **{k: V[k] for k in KEYS})
    d = np.random.default_rng(1).integers(0, 256, (500, 32), dtype=np.uint8)
    assert same_transform(G1.transform(d, 1), G2.transform(d, 1))
    with pytest.raises(OrbmError):
        ORBVocabulary.loadFromTextFile(matcher, str(tmp_path / "missing.txt"))
    (tmp_path / "bad.txt").write_text("99 3 0 0\n")
    with pytest.raises(OrbmError):
        ORBVocabulary.loadFromTextFile(matcher, str(tmp_path / "bad.txt"))
    bad = V["parent"].copy(); bad[3] = 9
    with pytest.raises(OrbmError):
        ORBVocabulary.from_arrays(matcher, V["k"], V["L"], 0, 0, bad, V["desc"], V["weight"], V["is_leaf"])
    leaf = V["is_leaf"].copy(); leaf[1] = 1
    with pytest.raises(OrbmError):
        ORBVocabulary.from_arrays(matcher, V["k"], V["L"], 0, 0, V["parent"], V["desc"], V["weight"], leaf)


def test_compute_bow_feeds_search_by_bow(B, oracle, matcher):
    """The chain the tracker runs: ComputeBoW on both frames, then SearchByBoW over the shared nodes.  Feature
    vectors are built on the device; the oracle gets the oracle's own transform."""
    from oracle import match_oracle as M
    from orb_slam_b200 import ORBmatcher, ORBVocabulary
    V = B.synth_vocabulary_fast(k=10, L=4, seed=21)
    O = B.Vocabulary.from_arrays(**{k: V[k] for k in KEYS})
    G = ORBVocabulary.from_arrays(matcher, **{k: V[k] for k in KEYS})
    pairs = [MD.make_pair(oracle, s, with_fv=False) for s in (0, 1)] + [MD.make_pair(oracle, 4, 333, 251, 700,
                                                                                      with_fv=False)]
    fr1 = [MD.with_valid(p[0], MD.random_valid(p[0].n, 0.7, 40 + k)) for k, p in enumerate(pairs)]
    fr2 = [MD.with_valid(p[1], MD.random_valid(p[1].n, 0.8, 50 + k)) for k, p in enumerate(pairs)]
    s1 = matcher.upload([MD.to_matchframe(f) for f in fr1]); s2 = matcher.upload([MD.to_matchframe(f) for f in fr2])
    s1.compute_bow(G, 2); s2.compute_bow(G, 2)
    for fs, fr in ((s1, fr1), (s2, fr2)):
        for k, f in enumerate(fr):
            (ow, ov), ofv = O.transform(f.desc, 2)
            f.featvec = ofv
            w, v = fs.bow_vector(k)
            assert np.array_equal(w, ow) and same_f64(v, ov)
            assert all(np.array_equal(x, y) for x, y in zip(fs.feature_vector(k), ofv))
    om = ORBmatcher(0.75, True, matcher=matcher)
    i1, i2 = [0, 1, 2, 0], [0, 1, 2, 1]
    for variant in ("kf_f", "kf_kf"):
        nm, rows = om.SearchByBoW(s1, s2, i1, i2, variant)
        for p, (a, b) in enumerate(zip(i1, i2)):
            on, orow = M.search_by_bow(fr1[a], fr2[b], variant, 0.75, True)
            assert nm[p] == on and np.array_equal(rows[p][: len(orow)], orow), (variant, p)
        assert nm[:3].sum() > 50
    sc = G.score(s1, s2, i1, i2)
    ref = [B.score_l1(O.transform(fr1[a].desc, 2)[0], O.transform(fr2[b].desc, 2)[0]) for a, b in zip(i1, i2)]
    assert same_f64(sc, ref) and sc[0] > sc[3]
    # recomputing with another levelsup replaces the vectors
    s1.compute_bow(G, 4)
    assert all(np.array_equal(x, y) for x, y in zip(s1.feature_vector(0), O.transform(fr1[0].desc, 4)[1]))


def test_extractor_to_bow_on_device(B, oracle, matcher):
    """E -> N1 -> N2 without the features leaving the GPU."""
    import torch
    from orb_slam_b200 import Camera, FrameSet, ORBextractor, ORBVocabulary
    H, W, NF = 240, 320, 500
    frames = np.stack([oracle.synth(H, W, s) for s in (3, 4, 5)])
    ex = ORBextractor(NF, 1.2, 6, 1, 20, max_width=W, max_height=H, max_batch=4)
    d_img = torch.from_numpy(frames).cuda()
    d_k = torch.zeros((3, NF, 28), dtype=torch.uint8, device="cuda")
    d_d = torch.zeros((3, NF, 32), dtype=torch.uint8, device="cuda")
    d_n = torch.zeros(3, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    ex.extract_batch_device(d_img.data_ptr(), 3, W, H, W, H * W, d_k.data_ptr(), d_d.data_ptr(), d_n.data_ptr(), st)
    cam = Camera(268.9633, 269.9858, 157.6087, 114.6369, 0.0, 0.0, 0.0, 0.0, 0.0, W, H)
    fs = FrameSet.from_extractor(matcher, d_k.data_ptr(), d_d.data_ptr(), d_n.data_ptr(), 3, NF, cam, st)
    V = B.synth_vocabulary_fast(k=10, L=3, seed=2)
    O = B.Vocabulary.from_arrays(**{k: V[k] for k in KEYS})
    G = ORBVocabulary.from_arrays(matcher, **{k: V[k] for k in KEYS})
    fs.compute_bow(G)           # levelsup = 4 >= L: everything files under the root, like a shallow ORBvoc would
    n = d_n.cpu().numpy(); hd = d_d.cpu().numpy()
    for f in range(3):
        (ow, ov), ofv = O.transform(hd[f, :n[f]], 4)
        w, v = fs.bow_vector(f)
        assert n[f] > 100 and np.array_equal(w, ow) and same_f64(v, ov)
        assert all(np.array_equal(x, y) for x, y in zip(fs.feature_vector(f), ofv))


def test_keyframe_database_candidate_scoring(matcher, B):
    """N4: orbm_bowdb_* (inverted file, shared-word counts, 0.8 * max cut, L1 scores of the survivors) against the oracle
    restatement of KeyFrameDatabase::DetectRelocalisationCandidates / DetectLoopCandidates (src/KeyFrameDatabase.cc:75-308),
    which tests/test_ref_matcher.py pins to the compiled reference: 60 keyframes, 9 queries, with and without excluded
    (covisible) keyframes; counts exact, scores compared as float bits."""
    from orb_slam_b200 import BowDatabase, MatchFrame, ORBVocabulary
    V = B.synth_vocabulary_fast(k=6, L=4, seed=3)
    ovoc = B.Vocabulary.from_arrays(**{k: V[k] for k in KEYS})
    voc = ORBVocabulary.from_arrays(matcher, **{k: V[k] for k in KEYS})
    rng = np.random.default_rng(8)
    leaves = V["desc"][V["is_leaf"] == 1]

    def frame(n, pool):
        d = leaves[rng.choice(pool, n)].copy()
        flip = rng.integers(0, 256, (n, 2))
        for c in range(2):
            d[np.arange(n), flip[:, c] >> 3] ^= (1 << (flip[:, c] & 7)).astype(np.uint8)
        z = np.zeros(n, np.float32)
        return MatchFrame(z, z, z, np.zeros(n, np.int32), d)
    pools = [rng.choice(len(leaves), 400, replace=False) for _ in range(6)]      # six "places" with their own words
    kfs = [frame(int(rng.integers(200, 900)), pools[k % 6]) for k in range(60)]
    kfs.append(frame(5, pools[0])); kfs.append(MatchFrame(*(np.zeros(0, np.float32),) * 3, np.zeros(0, np.int32), np.zeros((0, 32), np.uint8)))
    qs = [frame(int(rng.integers(300, 800)), pools[k % 6]) for k in range(9)]
    ks, qset = matcher.upload(kfs), matcher.upload(qs)
    ks.compute_bow(voc, 4); qset.compute_bow(voc, 4)
    kb = [ovoc.transform(f.desc, 4)[0] for f in kfs]; qb = [ovoc.transform(f.desc, 4)[0] for f in qs]
    db = BowDatabase(matcher, ks)
    order = [3, 0, 8, 1, 2, 4, 5, 6, 7, 3]                                        # any order, repeats allowed
    for exclude in (None, (rng.random((len(order), len(kfs))) < 0.2).astype(np.uint8)):
        common, mx, mn, score = db.query(qset, order, exclude)
        for r, q in enumerate(order):
            oc, omx, omn, osc = B.kfdb_candidates(qb[q], kb, None if exclude is None else exclude[r])
            assert np.array_equal(common[r], oc) and mx[r] == omx and mn[r] == omn, (r, q)
            assert np.array_equal(score[r].view(np.uint32), osc.view(np.uint32)), (r, q)
            assert (osc >= 0).sum() >= 3 and (osc < 0).sum() >= 30
    assert matcher.last_kernel_ms() > 0
    db.close()
