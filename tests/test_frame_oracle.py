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
