"""N1 pins: the oracle's restatement of cv::undistortPoints(src, dst, K, D, Mat(), K) — the call
Frame::UndistortKeyPoints / ComputeImageBounds make (reference src/Frame.cc:288-349) — is bit-exact against the
installed OpenCV (cv2 4.13), and the library's HOST-side image-bounds routine equals the oracle's."""
import numpy as np
import pytest

CAMS = [  # (fx, fy, cx, cy, k1, k2, p1, p2, k3, width, height)
    (268.9633, 269.9858, 157.6087, 114.6369, -0.4157, 0.2624, 0.0, 0.0, -0.1178, 320, 240),   # Data/Settings.yaml:6-16
    (517.3, 516.5, 318.6, 255.3, 0.2624, -0.9531, -0.0054, 0.0026, 1.1633, 640, 480),         # tangential terms
    (458.654, 457.296, 367.215, 248.375, -0.28340811, 0.07395907, 0.00019359, 1.76187114e-05, 0.0, 752, 480),
    (500.0, 500.0, 320.0, 240.0, 0.0, 0.0, 0.0, 0.0, 0.0, 640, 480),                           # k1 == 0: no undistortion
]


@pytest.mark.parametrize("cam", CAMS[:3])
def test_undistort_points_bit_exact_vs_cv2(cam):
    cv2 = pytest.importorskip("cv2")
    from oracle import frame_oracle as FO
    w, h = cam[9], cam[10]
    rng = np.random.default_rng(17)
    pts = np.concatenate([rng.uniform([-8, -8], [w + 8, h + 8], (100_000, 2)),
                          [[0, 0], [w, 0], [0, h], [w, h]]]).astype(np.float32)
    K = np.array([[cam[0], 0, cam[2]], [0, cam[1], cam[3]], [0, 0, 1]], np.float32)
    D = np.array(cam[4:9], np.float32).reshape(5, 1)
    ref = cv2.undistortPoints(pts.reshape(-1, 1, 2), K, D, None, K).reshape(-1, 2)
    got = FO.undistort_points(FO.Camera(*cam[:9]), pts)
    assert np.array_equal(ref.view(np.uint32), got.view(np.uint32))


@pytest.mark.parametrize("cam", CAMS)
def test_image_bounds_library_equals_oracle(cam):
    from oracle import frame_oracle as FO
    from orb_slam_b200 import Camera, image_bounds
    assert image_bounds(Camera(*cam)) == FO.image_bounds(FO.Camera(*cam[:9]), cam[9], cam[10])


def test_image_bounds_values_of_the_shipped_camera():
    """Known answer for Data/Settings.yaml on a 320x240 image: corners bulge outwards (barrel distortion removed)."""
    from oracle import frame_oracle as FO
    b, inv = FO.image_bounds(FO.Camera(*CAMS[0][:9]), 320, 240)
    assert b == (-64, 395, -43, 298)
    assert inv == (float(np.float32(64) / np.float32(459)), float(np.float32(48) / np.float32(341)))


# ---- N3: Frame::isInFrustum (src/Frame.cc:136-197) -------------------------------------------------------------------
def _cv2_is_in_frustum(cv2, Rcw, tcw, Ow, K, bounds, sf, pos, nrm, mn, mx, cos_limit):
    """The reference's statements with the REAL OpenCV doing the matrix expressions it has Python entry points for
    (cv2.gemm = `mRcw*P+mtcw`, cv2.norm = `cv::norm(PO)`); PO.dot(Pn) (no Python entry point) follows dotProd_'s
    scalar loop: double accumulation in element order."""
    f32 = np.float32
    fx, fy, cx, cy = K
    n = len(pos)
    iv = np.zeros(n, np.uint8); U = np.zeros(n, f32); V = np.zeros(n, f32); LV = np.full(n, -1, np.int32)
    VC = np.zeros(n, f32)
    for i in range(n):
        P = pos[i].reshape(3, 1)
        Pc = cv2.gemm(Rcw, P, 1.0, tcw.reshape(3, 1), 1.0)
        if Pc[2, 0] < 0.0:
            continue
        invz = f32(1.0 / np.float64(Pc[2, 0]))
        u = f32(f32(f32(fx * Pc[0, 0]) * invz) + cx)
        v = f32(f32(f32(fy * Pc[1, 0]) * invz) + cy)
        if u < f32(bounds[0]) or u > f32(bounds[1]) or v < f32(bounds[2]) or v > f32(bounds[3]):
            continue
        PO = (P - Ow.reshape(3, 1)).astype(f32)
        dist = f32(cv2.norm(PO))
        if dist < mn[i] or dist > mx[i]:
            continue
        dot = np.float64(0)
        for k in range(3):
            dot = dot + np.float64(PO[k, 0]) * np.float64(nrm[i, k])
        vc = f32(dot / np.float64(dist))
        if vc < f32(cos_limit):
            continue
        ratio = f32(dist / mn[i])
        lv = int(np.searchsorted(sf, ratio, side="left"))     # std::lower_bound
        iv[i] = 1; U[i] = u; V[i] = v; LV[i] = min(lv, len(sf) - 1); VC[i] = vc
    return iv, U, V, LV, VC


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_is_in_frustum_bit_exact_vs_cv2(seed):
    cv2 = pytest.importorskip("cv2")
    import matchdata as MD
    from oracle import frame_oracle as FO
    Rcw, tcw, K, bounds, sf, pos, nrm, mn, mx = MD.frustum_scene(seed, n=4000)
    view = FO.make_view(Rcw, tcw, *K, bounds, sf)
    Ow = np.array(view.Ow, np.float32)
    # the default Ow equals OpenCV's own evaluation of -mRcw.t()*mtcw
    assert np.array_equal(Ow, cv2.gemm(Rcw, tcw.reshape(3, 1), -1.0, None, 0.0, flags=cv2.GEMM_1_T).ravel())
    got = FO.is_in_frustum(view, pos, nrm, mn, mx, 0.5)
    ref = _cv2_is_in_frustum(cv2, Rcw, tcw, Ow, K, bounds, sf, pos, nrm, mn, mx, 0.5)
    assert np.array_equal(got[0], ref[0]) and np.array_equal(got[3], ref[3])
    for a, b in ((got[1], ref[1]), (got[2], ref[2]), (got[4], ref[4])):
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    # the scene takes every exit
    assert 0.05 < got[0].mean() < 0.6 and set(np.unique(got[3])) >= {-1, 1, 3, 7}   # level 0 needs dist == min_dist (boundary test below)


def test_is_in_frustum_boundaries():
    from oracle import frame_oracle as FO
    sf = np.cumprod(np.concatenate([[1.0], np.full(7, np.float32(1.2))]).astype(np.float32)).astype(np.float32)
    view = FO.make_view(np.eye(3), [0, 0, 0], 500.0, 500.0, 320.0, 240.0, (0, 640, 0, 480), sf)
    pos = np.array([[0, 0, 2], [0, 0, -1], [1.28, 0, 2], [1.2804, 0, 2], [0, 0, 2], [0, 0, 2], [0, 0, 2],
                    [0, 0, 2]], np.float32)
    nrm = np.array([[0, 0, 1]] * 6 + [[0, 0.8660254, 0.5], [0, 0.87, 0.49]], np.float32)
    mn = np.array([1, 1, 1, 1, 2, 0.1, 1, 1], np.float32)
    mx = np.array([4, 4, 4, 4, 4, 1.9, 4, 4], np.float32)
    iv, u, v, lv, vc = FO.is_in_frustum(view, pos, nrm, mn, mx, 0.5)
    #        centre  behind  u==640 kept  u>640  dist==min kept (ratio 1 -> level 0)  dist>max  cos==0.5 kept  cos<0.5
    assert list(iv) == [1, 0, 1, 0, 1, 0, 1, 0]
    assert u[0] == 320.0 and v[0] == 240.0 and u[2] == 640.0 and lv[4] == 0
    # ratio 2 -> first scale factor >= 2 is 1.2^4 = 2.0736 -> level 4
    assert lv[0] == 4 and vc[0] == 1.0
