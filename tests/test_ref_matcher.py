"""Matcher + Frame glue pins against the REFERENCE'S OWN CODE (Frame / KeyFrame / MapPoint / ORBmatcher /
ORBVocabulary compiled in place, oracle/ref_orbmatcher.cc; OpenCV routines from cv2).
 * golden (runs anywhere): oracle/match_oracle.cc reproduces tests/golden/ref_match.npz — grid CSR, GetFeaturesInArea,
   SearchForInitialization (matches and the updated vbPrevMatched), WindowSearch, SearchByBoW (KeyFrame-Frame and
   KeyFrame-KeyFrame).
 * live (authoring container): the whole chain image -> Frame (extract, undistort, bounds, grid) -> ComputeBoW ->
   searches, oracle against the compiled reference, plus MapPoint::ComputeDistinctiveDescriptors."""
import numpy as np
import pytest

import refgolden as RG
import refscene as RS


@pytest.fixture(scope="module")
def M(oracle):
    from oracle import match_oracle
    match_oracle.lib()
    return match_oracle


def test_oracle_reproduces_reference_golden(M):
    g = RG.load()
    for s in range(int(g["nscenes"])):
        fr = [RG.frame(g, s, 0), RG.frame(g, s, 1)]
        for f in (0, 1):
            p = f"s{s}_f{f}_"
            cs, it = M.grid_csr(fr[f])
            assert np.array_equal(cs, g[p + "cell_start"]) and np.array_equal(it, g[p + "grid_items"])
            st = g[p + "area_start"]
            for k, ((x, y, r), (l0, l1)) in enumerate(zip(g[p + "area_q"], g[p + "area_lv"])):
                got = M.features_in_area(fr[f], float(x), float(y), float(r), int(l0), int(l1))
                assert np.array_equal(got, g[p + "area_items"][st[k]:st[k + 1]]), (s, f, k)
        prev0 = np.stack([fr[0].x, fr[0].y], 1)
        for k, (win, ratio, ori) in enumerate(RG.SFI):
            prev = prev0.copy()
            n, m = M.search_for_initialization(fr[0], fr[1], prev, win, ratio, bool(ori))
            assert n == int(g[f"s{s}_sfi{k}_n"]) and np.array_equal(m, g[f"s{s}_sfi{k}_m"])
            assert np.array_equal(prev, g[f"s{s}_sfi{k}_prev"])
        for k, (win, l0, l1, ratio, ori) in enumerate(RG.WIN):
            n, m = M.window_search(fr[0], fr[1], win, ratio, bool(ori), l0, l1)
            assert n == int(g[f"s{s}_win{k}_n"]) and np.array_equal(m, g[f"s{s}_win{k}_m"])
        for k, (variant, ratio, ori) in enumerate(RG.BOW):
            n, m = M.search_by_bow(fr[0], fr[1], variant, ratio, bool(ori))
            assert n == int(g[f"s{s}_bow{k}_n"]) and np.array_equal(m, g[f"s{s}_bow{k}_m"])
            assert n > 50


def _rf():
    from oracle import ref_frontend as RF
    if not RF.available():
        pytest.skip("compiled reference front-end / cv2 not available here")
    return RF


@pytest.mark.parametrize("seed,cam,nfeat", [(21, RS.CAM_SHIPPED, 500), (22, RS.CAM_PINHOLE, 300)])
def test_chain_equals_compiled_reference(oracle, M, tmp_path, seed, cam, nfeat):
    RF = _rf()
    from oracle import bow_oracle as B
    from oracle import frame_oracle as FO
    V = B.synth_vocabulary(k=3, L=5, seed=seed, stop_frac=0.02)
    vpath = str(tmp_path / "voc.txt")
    B.save_text(vpath, **V)
    ovoc = B.Vocabulary.load_text(vpath)
    w = RF.World(vpath, nfeat, 1.2, 6, 1, 20)
    ims = RS.images(oracle, seed)
    E = oracle.OracleExtractor(nfeat, 1.2, 6, 1, 20, blur_mode=1, retain_mode=1)
    ocam = FO.Camera(*cam)
    ob, oinv = FO.image_bounds(ocam, 320, 240)
    rfs = [w.frame(ims[0], RS.K_of(cam), cam[4:9], reset_static=True), w.frame(ims[1], RS.K_of(cam), cam[4:9])]
    assert rfs[0].bounds == ob and rfs[0].inv == oinv                      # ComputeImageBounds, grid element sizes
    ofr = []
    for f, rf in enumerate(rfs):
        k, d = E(ims[f])                                                    # extraction
        assert np.array_equal(k, rf.kps) and np.array_equal(d, rf.desc)
        un = FO.undistort_points(ocam, np.stack([k["x"], k["y"]], 1)) if cam[4] != 0 else np.stack([k["x"], k["y"]], 1)
        assert np.array_equal(un[:, 0], rf.xu) and np.array_equal(un[:, 1], rf.yu)   # UndistortKeyPoints
        valid = RS.valid_flags(rf.n, seed + f)
        rf.set_valid(valid)
        (rw, rv), rfv = rf.compute_bow()                                    # Frame::ComputeBoW
        (ow, ov), ofv = ovoc.transform(d, 4)
        assert np.array_equal(rw, ow) and np.array_equal(rv.view(np.int64), ov.view(np.int64))
        assert all(np.array_equal(a, b) for a, b in zip(rfv, ofv)) and len(ofv[0]) > 1
        mf = RS.match_frame(rf, valid, ofv)
        ofr.append(mf)
        cs, it = rf.grid()
        ocs, oit = M.grid_csr(mf)
        assert np.array_equal(cs, ocs) and np.array_equal(it, oit)          # grid binning
    prev0 = np.stack([ofr[0].x, ofr[0].y], 1)
    for win, ratio, ori in ((100, 0.9, True), (30, 0.8, False)):
        rn, rm, rp = RF.search_for_initialization(rfs[0], rfs[1], prev0, win, ratio, ori)
        op = prev0.copy()
        on, om = M.search_for_initialization(ofr[0], ofr[1], op, win, ratio, ori)
        assert rn == on and np.array_equal(rm, om) and np.array_equal(rp, op) and on > 30
    for win, l0, l1, ratio, ori in ((200, 0, 2**31 - 1, 0.8, True), (100, 1, 3, 0.8, False), (50, -1, 2**31 - 1, 0.9, True)):
        rn, rm = RF.window_search(rfs[0], rfs[1], win, l0, l1, ratio, ori)
        on, om = M.window_search(ofr[0], ofr[1], win, ratio, ori, l0, l1)
        assert rn == on and np.array_equal(rm, om)
    for variant in ("kf_f", "kf_kf"):
        for ratio, ori in ((0.75, True), (0.6, False)):
            rn, rm = RF.search_by_bow(rfs[0], rfs[1], variant, ratio, ori)
            on, om = M.search_by_bow(ofr[0], ofr[1], variant, ratio, ori)
            assert rn == on and np.array_equal(rm, om) and on > 30


def test_oracle_reproduces_reference_projection_golden(M):
    """Runs anywhere: the oracle against tests/golden/ref_proj.npz — the reference's own SearchByProjection overloads
    (local map, last frame, relocalisation, Sim3; tests/golden/make_ref_proj_golden.py)."""
    names = set()
    for c in RG.proj_cases():
        on, om = M.search_by_projection(c["target"], c["points"], c["variant"], c["th_dist"], c["ratio"], c["ori"])
        assert on == c["ref_n"] and np.array_equal(om, c["ref_m"]), c["name"]
        names.add(c["name"].split(":")[1])
    assert names == {"local_map", "last_frame", "reloc", "sim3"}


def test_distinctive_descriptors_equal_compiled_reference(oracle, M):
    RF = _rf()
    w = RF.World(None, 300, 1.2, 4, 1, 20)
    cam = RS.CAM_PINHOLE
    frames = []
    for s in range(6):
        im, _ = RS.images(oracle, 40 + s)
        f = w.frame(im, RS.K_of(cam), cam[4:9], reset_static=(s == 0))
        f.set_valid(np.zeros(f.n, np.uint8))
        frames.append(f)
    kfs = [f.keyframe() for f in frames]
    rng = np.random.default_rng(2)
    for nobs in (1, 2, 3, 5, 6):
        for _ in range(6):
            sel = rng.choice(6, nobs, replace=False)
            idx = [int(rng.integers(0, frames[k].n)) for k in sel]
            got, order = w.distinctive([kfs[k] for k in sel], idx)
            # descriptors in the order the reference's std::map<KeyFrame*, size_t> visits the observations
            desc = np.stack([frames[sel[o]].desc[idx[o]] for o in order])
            best = M.distinctive_descriptors(desc, [0, nobs])[0]
            assert np.array_equal(got, desc[best])


@pytest.mark.parametrize("seed,cam", [(31, RS.CAM_SHIPPED), (32, RS.CAM_PINHOLE), (71, RS.CAM_SHIPPED)])
def test_search_by_projection_equals_compiled_reference(oracle, M, seed, cam):
    """M6, all four overloads run by the reference on its own Frame / KeyFrame / MapPoint objects (scenes and the
    caller-side float expressions: tests/refscene.py::projection_cases) against the oracle's flat-array form:
    local map SearchByProjection(Frame, vector<MapPoint*>, th) (src/ORBmatcher.cc:49-125), last frame
    SearchByProjection(CurrentFrame, LastFrame, th) (:1507-1620), relocalisation SearchByProjection(CurrentFrame, pKF,
    sAlreadyFound, th, ORBdist) (:1622-1746), loop closing SearchByProjection(pKF, Scw, vpPoints, vpMatched, th)
    (:286-407); poses at identity, so the pose algebra returns the world point unchanged and everything else is the
    reference's own code."""
    RF = _rf()
    seen = set()
    for c in RS.projection_cases(oracle, RF, seed, cam):
        on, om = M.search_by_projection(c["target"], c["points"], c["variant"], c["th_dist"], c["ratio"], c["ori"])
        assert c["ref_n"] == on and np.array_equal(c["ref_m"], om) and on > 15, (c["name"], c["ref_n"], on)
        seen.add(c["name"].split(":")[0])
    assert seen == {"local_map", "last_frame", "reloc", "sim3"}


@pytest.mark.parametrize("seed,cam", [(81, RS.CAM_SHIPPED), (82, RS.CAM_PINHOLE)])
def test_is_in_frustum_equals_compiled_reference(oracle, seed, cam):
    """N3: Frame::isInFrustum (src/Frame.cc:136-197) run by the reference's own Frame on its own MapPoints (pose at
    identity through UpdatePoseMatrices, so mRcw*P+mtcw and P-mOw return P) against the oracle: return value,
    mbTrackInView, mTrackProjX/Y, mnTrackScaleLevel, mTrackViewCos compared as raw bits.  The matrix expressions under a
    general pose are pinned against OpenCV itself in tests/test_frame_oracle.py."""
    RF = _rf()
    from oracle import frame_oracle as FO
    w = RF.World(None, 400, 1.2, 6, 1, 20)
    ims = RS.images(oracle, seed)
    f1 = w.frame(ims[0], RS.K_of(cam), cam[4:9], reset_static=True)
    f2 = w.frame(ims[1], RS.K_of(cam), cam[4:9])
    sf = np.cumprod(np.concatenate([[1.0], np.full(5, np.float32(1.2))]).astype(np.float32)).astype(np.float32)
    rng = np.random.default_rng(seed)
    n = f1.n
    fx, fy, cx, cy = (np.float32(c) for c in cam[:4])
    z = rng.uniform(0.5, 9.0, n).astype(np.float32)
    z[rng.random(n) < 0.08] *= -1                                         # behind the camera
    tu = rng.uniform(-60, 380, n); tv = rng.uniform(-60, 300, n)          # a band outside the image on every side
    pos = np.stack([(tu - cx) / fx * z, (tv - cy) / fy * z, z], 1).astype(np.float32)
    dist = RS._norm3(pos)
    min_dist = (dist / rng.uniform(0.7, 1.2 ** 7, n).astype(np.float32)).astype(np.float32)    # some closer than min
    max_dist = (min_dist * rng.choice(np.array([1.5, 3.0, 1.2 ** 6], np.float32), n)).astype(np.float32)
    d = rng.normal(0, 1, (n, 3)); d /= np.linalg.norm(d, axis=1)[:, None]
    normal = (0.6 * pos / np.abs(dist[:, None]) + 0.8 * d).astype(np.float32)   # cosines spread around the 0.5 limit
    f1.set_valid(np.ones(n, np.uint8))
    f1.set_point_geometry(pos, min_dist, max_dist, normal)
    view = FO.make_view(np.eye(3), [0, 0, 0], fx, fy, cx, cy, (f2.bounds[0], f2.bounds[1], f2.bounds[2], f2.bounds[3]), sf)
    for lim in (0.5, 0.8):
        ret, iv, u, v, lv, vc = RF.is_in_frustum(f2, f1, lim)
        o_iv, o_u, o_v, o_lv, o_vc = FO.is_in_frustum(view, pos, normal, min_dist, max_dist, lim)
        assert np.array_equal(ret, iv) and np.array_equal(iv, o_iv)
        k = iv == 1
        assert 0.03 < k.mean() < 0.7 and len(set(lv[k].tolist())) >= 5
        assert np.array_equal(lv[k], o_lv[k])
        for a, b in ((u, o_u), (v, o_v), (vc, o_vc)):
            assert np.array_equal(a[k].view(np.uint32), b[k].view(np.uint32))


@pytest.mark.parametrize("ori", [True, False])
def test_search_for_triangulation_equals_compiled_reference(oracle, M, tmp_path, ori):
    """M9: SearchForTriangulation (src/ORBmatcher.cc:852-1014, CheckDistEpipolarLine :136-153) between two real
    KeyFrames (feature vectors from the reference's ComputeBoW, MapPoints on some keypoints)."""
    RF = _rf()
    from oracle import bow_oracle as B
    V = B.synth_vocabulary(k=3, L=5, seed=5, stop_frac=0.0)
    vpath = str(tmp_path / "voc.txt")
    B.save_text(vpath, **V)
    cam = RS.CAM_SHIPPED
    w = RF.World(vpath, 500, 1.2, 6, 1, 20)
    ims = RS.images(oracle, 51, shift=(9, -7))
    rfs = [w.frame(ims[0], RS.K_of(cam), cam[4:9], reset_static=True), w.frame(ims[1], RS.K_of(cam), cam[4:9])]
    ofr = []
    for f, rf in enumerate(rfs):
        valid = (np.random.default_rng(60 + f).random(rf.n) < 0.3).astype(np.uint8)   # keypoints that already have a MapPoint
        rf.set_valid(valid)
        _, fv = rf.compute_bow()
        ofr.append(RS.match_frame(rf, valid, fv))
    sf = np.cumprod(np.concatenate([[1.0], np.full(5, np.float32(1.2))]).astype(np.float32)).astype(np.float32)
    sigma2 = (sf * sf).astype(np.float32)                                     # mvLevelSigma2 (src/Frame.cc:99-103)
    t = np.array([9.0, -7.0, 0.0]) + 1e-3
    F12 = np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]], np.float32) * np.float32(0.05)
    for ratio in (0.6, 0.8):
        rn, rm = RF.search_for_triangulation(rfs[0], rfs[1], F12, ratio, ori)
        on, om = M.search_for_triangulation(ofr[0], ofr[1], F12, sigma2, ori)
        assert rn == on and np.array_equal(rm, om) and on > 20, (ratio, ori)


@pytest.mark.parametrize("seed,cam", [(41, RS.CAM_SHIPPED), (42, RS.CAM_PINHOLE)])
def test_fuse_sim3_projection_frames_equal_compiled_reference(oracle, M, seed, cam):
    """The rest of M6 and M4: ORBmatcher::Fuse (src/ORBmatcher.cc:1016-1134 and :1136-1265), SearchBySim3 (:1267-1505) and
    SearchByProjection(F1, F2, window) (:519-594) run by the compiled reference on its own KeyFrame / MapPoint objects,
    against the oracle's independent best-only scoring.  Fuse returns only a count; the per-point keypoints are read
    off what it does to the object graph (observations added, points / keyframe points replaced), which
    refscene.fuse_expectation predicts from the oracle's per-point result."""
    RF = _rf()
    seen = {}
    for c in RS.fuse_cases(oracle, RF, seed, cam):
        kind = c["kind"]
        seen[kind] = seen.get(kind, 0) + 1
        if kind in ("fuse", "fuse_sim3"):
            n, best = M.search_by_projection_independent(c["target"], c["points"], c["th_dist"])
            assert n == c["ref_n"] and n > 40, (c["name"], n, c["ref_n"])
            exp = RS.fuse_expectation(c, best)
            for key, val in exp.items():
                assert np.array_equal(val, c[key]), (c["name"], key)
            tv = c["tvalid"][best[best >= 0]]
            assert (tv == 0).sum() > 5 and (tv == 1).sum() > 5           # both branches taken
            assert len(np.unique(best[best >= 0])) < (best >= 0).sum()   # two points chose one keypoint
        elif kind == "sim3":
            n, m12 = M.search_by_sim3(c["f1"], c["f2"], c["p12"], c["p21"], c["th_dist"])
            exp = np.where(c["pre"] >= 0, c["pre"], m12)                 # pre-set entries stay (vpMatches12 in/out)
            assert n == c["ref_n"] and np.array_equal(exp, c["ref_m"]) and n > 15, (c["name"], n, c["ref_n"])
        else:
            n, m21 = M.search_by_projection_ff(c["f1"], c["f2"], c["uv"], c["window"], c["ratio"])
            assert n == c["ref_n"] and np.array_equal(m21, c["ref_m"]) and n > 15, (c["name"], n, c["ref_n"])
    assert seen == {"fuse": 2, "fuse_sim3": 2, "sim3": 2, "proj_ff": 2}


def test_oracle_reproduces_reference_fuse_golden(M):
    """Runs anywhere (no compiled reference needed): the oracle against tests/golden/ref_fuse.npz."""
    kinds = {}
    for c in RG.fuse_cases():
        kinds[c["kind"]] = kinds.get(c["kind"], 0) + 1
        if c["kind"] in ("fuse", "fuse_sim3"):
            n, best = M.search_by_projection_independent(c["target"], c["points"], c["th_dist"])
            assert n == c["ref_n"], c["name"]
            for key, val in RS.fuse_expectation(c, best).items():
                assert np.array_equal(val, c[key]), (c["name"], key)
        elif c["kind"] == "sim3":
            n, m12 = M.search_by_sim3(c["f1"], c["f2"], c["p12"], c["p21"], c["th_dist"])
            assert n == c["ref_n"] and np.array_equal(np.where(c["pre"] >= 0, c["pre"], m12), c["ref_m"]), c["name"]
        else:
            n, m21 = M.search_by_projection_ff(c["f1"], c["f2"], c["uv"], c["window"], c["ratio"])
            assert n == c["ref_n"] and np.array_equal(m21, c["ref_m"]), c["name"]
    assert kinds == {"fuse": 4, "fuse_sim3": 4, "sim3": 4, "proj_ff": 4}


def _kfdb_scene(oracle, RF, tmp_path, seed, nkf=7):
    """nkf keyframes + one query view of synthetic scenes (two scenes, several shifts) with a 5^4-word vocabulary."""
    from oracle import bow_oracle as B
    V = B.synth_vocabulary(k=5, L=4, seed=seed, stop_frac=0.0)
    vpath = str(tmp_path / "voc.txt")
    B.save_text(vpath, **V)
    w = RF.World(vpath, 300, 1.2, 6, 1, 20)
    cam = RS.CAM_PINHOLE
    frames = []
    for k in range(nkf + 1):
        a, b = RS.images(oracle, seed + (k % 3), shift=(3 * k, 2 * k))
        f = w.frame(b if k % 2 else a, RS.K_of(cam), cam[4:9], reset_static=(k == 0))
        frames.append(f)
    bows = [f.compute_bow()[0] for f in frames]                # Frame::ComputeBoW -> (words, weights)
    return w, frames, bows


def test_keyframe_database_candidate_scoring_equals_compiled_reference(oracle, tmp_path):
    """N4: the compiled KeyFrameDatabase (add, DetectRelocalisationCandidates :198-308, DetectLoopCandidates :75-196) on
    real KeyFrames against oracle/bow_oracle.py::kfdb_candidates: shared-word counts, the 0.8 * max cut, float scores;
    without covisibility links the returned candidates are the keyframes whose score exceeds 0.75 * the best score."""
    RF = _rf()
    from oracle import bow_oracle as B
    w, frames, bows = _kfdb_scene(oracle, RF, tmp_path, 51)
    kfs, query = frames[:-1], frames[-1]
    RF.kfdb_add(w, kfs)
    n, touched, words, score, cand = RF.detect_relocalisation_candidates(w, query, kfs)
    common, mx, mn, osc = B.kfdb_candidates(bows[-1], bows[:-1])
    assert np.array_equal(touched.astype(bool), common > 0) and np.array_equal(words[common > 0], common[common > 0])
    above = common > mn
    assert above.sum() >= 2 and (~above).sum() >= 1 and mx > 20
    assert np.array_equal(score[above].view(np.uint32), osc[above].view(np.uint32))          # float bits
    best = osc[above].max()
    assert np.array_equal(cand.astype(bool), above & (osc > np.float32(0.75) * best)) and n == int(cand.sum())
    # loop detection: the query is keyframe 2 (mnId != 0: keyframes start with mnLoopQuery == 0), keyframes 1 and 3 are its
    # covisibility neighbours (skipped, :95)
    connected = np.zeros(len(kfs), np.uint8); connected[[1, 3]] = 1
    connected[2] = 1                                            # the query itself shares every word with itself
    n, touched, words, score, cand = RF.detect_loop_candidates(w, kfs[2], kfs, connected, 0.0)
    common, mx, mn, osc = B.kfdb_candidates(bows[2], bows[:-1], exclude=connected)
    free = connected == 0
    assert np.array_equal(touched[free].astype(bool), common[free] > 0) and not touched[~free].any()
    assert np.array_equal(words[free & (common > 0)], common[free & (common > 0)])
    above = free & (common > mn)
    assert above.sum() >= 1
    assert np.array_equal(score[above].view(np.uint32), osc[above].view(np.uint32))
