"""Extractor pin against the REFERENCE'S OWN CODE: /root/reference/src/ORBextractor.cc compiled in place with its OpenCV
library calls served by the real OpenCV (cv2 4.13) — see oracle/ref_extractor.py, oracle/ref_shim_cv/.
 * golden: tests/golden/ref_extract.npz was produced by that build (tests/golden/make_ref_extract_golden.py); the
   restatement oracle/orb_oracle.cc must reproduce it bit for bit (blur_mode = 1: OpenCV >= 3.4 Gaussian taps;
   retain_mode = the retainBest variant the vectors were made with).  Runs anywhere.
 * live (authoring container only: needs oracle/_ref/libref_orbextractor.so + cv2): more frames, both headline
   configurations, compared field by field."""
import os

import numpy as np
import pytest

FIELDS = ("x", "y", "size", "angle", "response", "octave", "class_id")
GOLD = os.path.join(os.path.dirname(__file__), "golden", "ref_extract.npz")


def same_keypoints(a, b):
    return len(a) == len(b) and all(np.array_equal(a[f].view(np.uint32) if a[f].dtype.kind == "f" else a[f],
                                                   b[f].view(np.uint32) if b[f].dtype.kind == "f" else b[f])
                                    for f in FIELDS)


def test_oracle_reproduces_reference_golden(oracle):
    g = np.load(GOLD)
    for c in range(int(g["ncases"])):
        nf, nl, st, th, sf = (int(v) for v in g[f"params{c}"])
        for mode in (1, 0):
            E = oracle.OracleExtractor(nf, sf / 1000.0, nl, st, th, blur_mode=1, retain_mode=mode)
            k, d = E(g[f"img{c}"])
            assert same_keypoints(k, g[f"kps{c}_{mode}"]), (c, mode)
            assert np.array_equal(d, g[f"desc{c}_{mode}"]), (c, mode)
    assert len(g["kps3_1"]) == 500 and len(g["kps4_1"]) > 0


def _ref():
    from oracle import ref_extractor as RE
    if not RE.available():
        pytest.skip("compiled reference extractor / cv2 not available here")
    return RE


@pytest.mark.parametrize("cfg", [(480, 640, 1000, 8, 1, 20, (0, 1, 2)),        # BASELINE configs 1 / 2
                                 (376, 1241, 2000, 8, 1, 20, (0, 1)),          # config 3
                                 (480, 640, 1000, 8, 0, 20, (3,)),             # Harris score
                                 (243, 317, 700, 5, 1, 35, (4,))])
def test_oracle_equals_compiled_reference(oracle, cfg):
    RE = _ref()
    h, w, nf, nl, st, th, seeds = cfg
    R = RE.RefExtractor(nf, 1.2, nl, st, th)
    for mode in (1, 0):
        RE.set_retain_mode(mode)
        E = oracle.OracleExtractor(nf, 1.2, nl, st, th, blur_mode=1, retain_mode=mode)
        for s in seeds:
            img = oracle.synth(h, w, s)
            rk, rd = R(img)
            ok, od = E(img)
            assert same_keypoints(ok, rk) and np.array_equal(od, rd), (cfg, mode, s)
            assert len(rk) > nf // 2
    RE.set_retain_mode(1)


def test_compiled_reference_edge_images(oracle):
    RE = _ref()
    R = RE.RefExtractor(300, 1.2, 4, 1, 20)
    E = oracle.OracleExtractor(300, 1.2, 4, 1, 20, blur_mode=1, retain_mode=1)
    flat = np.full((120, 160), 77, np.uint8)
    assert len(R(flat)[0]) == 0 and len(E(flat)[0]) == 0
    low = (oracle.synth(120, 160, 9).astype(np.int32) // 8 + 90).astype(np.uint8)   # the threshold-7 retry everywhere
    rk, rd = R(low)
    ok, od = E(low)
    assert same_keypoints(ok, rk) and np.array_equal(od, rd)


@pytest.mark.parametrize("cfg", [(480, 640, 1000, 8, 1, 20, (0, 1, 2)), (376, 1241, 2000, 8, 1, 20, (0,)),
                                 (480, 640, 1000, 8, 0, 20, (3,)), (243, 317, 700, 5, 1, 35, (4,))])
def test_native_bridged_reference_equals_oracle_and_cv2_bridged(oracle, cfg):
    """The timed CPU baseline (bench.py cpu_baseline.kind = "reference"): the SAME compiled src/ORBextractor.cc with its
    OpenCV routines served natively (oracle/ref_native_bridge.cc) returns exactly what the oracle returns in every
    OpenCV-era mode, and, where cv2 is importable, exactly what the cv2-bridged build returns."""
    from oracle import ref_extractor as RE
    if not RE.native_available():
        pytest.skip("oracle/_ref/libref_orbextractor_native.so not built here")
    h, w, nf, nl, st, th, seeds = cfg
    imgs = np.stack([oracle.synth(h, w, s) for s in seeds])
    for retain, blur in ((1, 1), (0, 0), (0, 1)):
        N = RE.NativeRefExtractor(nf, 1.2, nl, st, th, retain_mode=retain, blur_mode=blur)
        E = oracle.OracleExtractor(nf, 1.2, nl, st, th, blur_mode=blur, retain_mode=retain)
        k, d, n = N.extract_many(imgs, nthreads=min(len(seeds), 3))
        for i in range(len(seeds)):
            ok, od = E(imgs[i])
            assert same_keypoints(ok, k[i, :n[i]]) and np.array_equal(od, d[i, :n[i]]), (cfg, retain, blur, i)
    if RE.available():
        R = RE.RefExtractor(nf, 1.2, nl, st, th)
        RE.set_retain_mode(1)
        N = RE.NativeRefExtractor(nf, 1.2, nl, st, th, retain_mode=1, blur_mode=1)
        rk, rd = R(imgs[0])
        nk, nd = N(imgs[0])
        assert same_keypoints(nk, rk) and np.array_equal(nd, rd)
