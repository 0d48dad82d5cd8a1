"""Pins the oracle's OpenCV primitive models bit-for-bit against the real cv2 (4.13) primitives — 'Oracle B' of
SURVEY.md §7.1 step 1.  The reference ships no tests; this is what anchors the restatement."""
import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")


def _imgs(oracle):
    rng = np.random.default_rng(3)
    return [oracle.synth(480, 640, 0), oracle.synth(376, 1241, 5), rng.integers(0, 256, (97, 131), dtype=np.uint8)]


def test_resize_chain(oracle):
    for img in _imgs(oracle):
        cur = img
        inv = np.float32(1.0)
        for _ in range(7):
            inv = np.float32(inv * np.float32(1.0 / np.float64(np.float32(1.2))))
            w = int(np.rint(np.float32(img.shape[1]) * inv)); h = int(np.rint(np.float32(img.shape[0]) * inv))
            if w < 8 or h < 8:
                break
            a = oracle.resize(cur, w, h)
            b = cv2.resize(cur, (w, h), interpolation=cv2.INTER_LINEAR)
            assert np.array_equal(a, b)
            cur = a


def test_fast_nms_cells(oracle):
    for img in _imgs(oracle):
        for th in (20, 7):
            det = cv2.FastFeatureDetector_create(th, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
            H, W = img.shape
            for (x0, y0, w, h) in [(0, 0, W, H), (13, 13, min(128, W - 13), min(81, H - 13)), (40, 30, 55, 32),
                                   (5, 5, 7, 7), (5, 5, 6, 20)]:
                sub = img[y0:y0 + h, x0:x0 + w]
                kp = det.detect(np.ascontiguousarray(sub))
                xs, ys, sc = oracle.fast(sub, th)
                assert len(kp) == len(xs)
                got = [(int(k.pt[0]), int(k.pt[1]), int(k.response)) for k in kp]
                assert got == list(zip(xs.tolist(), ys.tolist(), sc.tolist()))  # same order too
                assert all(k.size == 7 and k.angle == -1 and k.octave == 0 and k.class_id == -1 for k in kp)


def test_threshold_free_score_map(oracle):
    """kept(t=20) == {k in kept(t=7): S >= 20}: one score map serves the fastTh pass and the th=7 retry."""
    img = oracle.synth(200, 300, 11)
    x7, y7, s7 = oracle.fast(img, 7)
    x20, y20, s20 = oracle.fast(img, 20)
    m = s7 >= 20
    assert np.array_equal(x7[m], x20) and np.array_equal(y7[m], y20) and np.array_equal(s7[m], s20)


def test_blur_fixed_point_mode(oracle):
    for img in _imgs(oracle):
        a = oracle.blur7(img, 1)
        b = cv2.GaussianBlur(img, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
        assert np.array_equal(a, b)


def test_blur_cv24_mode(oracle):
    """OpenCV 2.4 8-bit path = taps cvRound(k*256).  cv2 4.13's sepFilter2D evaluates the same taps in float with
    round-half-even, so the two agree except on exact .5 ties of (sum+32768)>>16."""
    k = cv2.getGaussianKernel(7, 2, cv2.CV_32F).ravel()
    taps = np.rint(k * 256).astype(int)
    assert list(taps) == [18, 34, 49, 55, 49, 34, 18]
    kq = (taps / 256.0).astype(np.float32).reshape(7, 1)
    for img in _imgs(oracle):
        a = oracle.blur7(img, 0)
        b = cv2.sepFilter2D(img, -1, kq, kq, borderType=cv2.BORDER_REFLECT_101)
        p = np.pad(img.astype(np.int64), 3, mode="reflect")
        H, W = img.shape
        hs = sum(int(taps[i]) * p[:, i:i + W] for i in range(7))
        vs = sum(int(taps[j]) * hs[j:j + H, :] for j in range(7))
        tie = (vs & 0xFFFF) == 0x8000
        assert np.array_equal(a[~tie], b[~tie])
        assert np.array_equal(a, np.minimum((vs + 32768) >> 16, 255).astype(np.uint8))


def test_fast_atan2(oracle):
    rng = np.random.default_rng(1)
    v = rng.integers(-300000, 300000, (20000, 2)).astype(np.float32)
    for y, x in v:
        assert np.float32(cv2.fastAtan2(float(y), float(x))) == np.float32(oracle.fast_atan2(float(y), float(x)))


def test_copy_make_border_reflect(oracle):
    """Oracle level images inside the extractor equal cv2.resize chains; border = copyMakeBorder(REFLECT_101)."""
    img = oracle.synth(120, 160, 2)
    E = oracle.OracleExtractor(300, 1.2, 4, 1, 20)
    E(img)
    cur = img
    for l in range(4):
        w, h = E.level_size(l)
        if l:
            cur = cv2.resize(cur, (w, h), interpolation=cv2.INTER_LINEAR)
        assert np.array_equal(E.level_image(l, 0), cur)
