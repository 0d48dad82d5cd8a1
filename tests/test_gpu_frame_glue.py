"""N1: orbm_frameset_from_extractor == the reference's Frame constructor glue (src/Frame.cc:59-127) done by
the oracle on the host: undistorted coordinates bit-exact, SoA columns, 64x48 grid CSR, and a search on the
device-built set equal to the same search on a host-uploaded set."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CAMS = [
    (268.9633 * 2, 269.9858 * 2, 157.6087 * 2, 114.6369 * 2, -0.4157, 0.2624, 0.0, 0.0, -0.1178, 640, 480),
    (517.3, 516.5, 318.6, 255.3, 0.2624, -0.9531, -0.0054, 0.0026, 1.1633, 640, 480),
    (500.0, 500.0, 320.0, 240.0, 0.0, 0.0, 0.0, 0.0, 0.0, 640, 480),
]


@pytest.mark.parametrize("cam", CAMS)
def test_frameset_from_extractor(oracle, cam):
    import torch
    from orb_slam_b200 import (Camera, FrameSet, KP_DTYPE, Matcher, MatchFrame, ORBextractor, ORBmatcher)
    from oracle import frame_oracle as FO
    from oracle import match_oracle as MO
    F, H, W, NF = 4, 480, 640, 800
    frames = np.stack([oracle.synth(H, W, 900 + s) for s in range(F)])
    frames[3] = 0                                            # a frame without keypoints
    ex = ORBextractor(NF, 1.2, 8, 1, 20, max_width=W, max_height=H, max_batch=F)
    d_img = torch.from_numpy(frames).cuda()
    d_k = torch.zeros((F, NF, 28), dtype=torch.uint8, device="cuda")
    d_d = torch.zeros((F, NF, 32), dtype=torch.uint8, device="cuda")
    d_n = torch.zeros(F, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    ex.extract_batch_device(d_img.data_ptr(), F, W, H, W, H * W, d_k.data_ptr(), d_d.data_ptr(), d_n.data_ptr(), st)
    m = Matcher(0)
    camera = Camera(*cam)
    fs = FrameSet.from_extractor(m, d_k.data_ptr(), d_d.data_ptr(), d_n.data_ptr(), F, NF, camera, st)
    n = d_n.cpu().numpy()
    assert np.array_equal(fs.counts, n) and n[3] == 0 and n[0] > 100
    kps = d_k.cpu().numpy().view(KP_DTYPE).reshape(F, NF)
    desc = d_d.cpu().numpy()
    ocam = FO.Camera(*cam[:9])
    (mnx, mxx, mny, mxy), _ = FO.image_bounds(ocam, W, H)
    host_frames = []
    for f in range(F):
        k = kps[f, :n[f]]
        xy = np.stack([k["x"], k["y"]], 1).astype(np.float32)
        un = FO.undistort_points(ocam, xy) if cam[4] != 0 else xy
        x, y, a, o, d = fs.keypoints(f)
        assert np.array_equal(x.view(np.uint32), un[:, 0].copy().view(np.uint32)), f"x frame {f}"
        assert np.array_equal(y.view(np.uint32), un[:, 1].copy().view(np.uint32)), f"y frame {f}"
        assert np.array_equal(a.view(np.uint32), k["angle"].view(np.uint32)) and np.array_equal(o, k["octave"])
        assert np.array_equal(d, desc[f, :n[f]])
        mf = MatchFrame(un[:, 0], un[:, 1], k["angle"], k["octave"], desc[f, :n[f]], min_x=float(mnx), min_y=float(mny),
                        max_x=float(mxx), max_y=float(mxy))
        host_frames.append(mf)
        cs, items = fs.grid(f)
        ocs, oitems = MO.grid_csr(mf)
        assert np.array_equal(cs, ocs) and np.array_equal(items, oitems), f"grid frame {f}"
    # the device-built set behaves like a host-uploaded one in a search (frame 0 -> 1, 1 -> 2, 2 -> 0)
    om = ORBmatcher(0.9, True, matcher=m)
    hs = om.upload(host_frames)
    i1, i2 = np.array([0, 1, 2], np.int32), np.array([1, 2, 0], np.int32)
    prev_a = [np.stack([host_frames[i].x, host_frames[i].y], 1).copy() for i in i1]
    prev_b = [p.copy() for p in prev_a]
    nm_a, rows_a = om.SearchForInitialization(fs, fs, i1, i2, prev_a, 100)
    nm_b, rows_b = om.SearchForInitialization(hs, hs, i1, i2, prev_b, 100)
    assert np.array_equal(nm_a, nm_b)
    for p in range(3):
        assert np.array_equal(rows_a[p, :n[i1[p]]], rows_b[p, :n[i1[p]]])
        assert np.array_equal(prev_a[p], prev_b[p])
    fs.close(); hs.close(); ex.close()


def test_is_in_frustum_equals_oracle():
    """N3: batched Frame::isInFrustum (several views x the same map points) — flags, levels and raw float bits."""
    import matchdata as MD
    from oracle import frame_oracle as FO
    from orb_slam_b200 import FrustumView, Matcher
    m = Matcher(0)
    scenes = [MD.frustum_scene(s, n=20000) for s in (0, 1, 2)]
    pos, nrm, mn, mx = scenes[0][5:]
    views, oviews = [], []
    for Rcw, tcw, K, bounds, sf, *_ in scenes:
        ov = FO.make_view(Rcw, tcw, *K, bounds, sf)
        oviews.append(ov)
        views.append(FrustumView.make(Rcw, tcw, np.array(ov.Ow, np.float32), *K, bounds, sf))
    # a 5-level pyramid as well (clamp to nlevels - 1)
    ov = FO.make_view(scenes[0][0], scenes[0][1], *scenes[0][2], scenes[0][3], scenes[0][4][:5])
    oviews.append(ov)
    views.append(FrustumView.make(scenes[0][0], scenes[0][1], np.array(ov.Ow, np.float32), *scenes[0][2], scenes[0][3],
                                  scenes[0][4][:5]))
    iv, u, v, lv, vc = m.is_in_frustum(views, pos, nrm, mn, mx, 0.5)
    for w, ov in enumerate(oviews):
        oiv, ou, ovv, olv, ovc = FO.is_in_frustum(ov, pos, nrm, mn, mx, 0.5)
        assert np.array_equal(iv[w], oiv) and np.array_equal(lv[w], olv), w
        for a, b in ((u[w], ou), (v[w], ovv), (vc[w], ovc)):
            assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), w
    assert iv[0].mean() > 0.1 and lv[3].max() == 4
    # the accepted points feed SearchByProjection's point records
    from orb_slam_b200 import ProjectedPoints
    ok = iv[0] == 1
    pts = ProjectedPoints.for_local_map(u[0][ok], v[0][ok], vc[0][ok], lv[0][ok],
                                        np.zeros((int(ok.sum()), 32), np.uint8), scenes[0][4])
    assert pts.n == int(ok.sum())
    e = m.is_in_frustum(views[:1], np.zeros((0, 3), np.float32), np.zeros((0, 3), np.float32), [], [])
    assert e[0].shape == (1, 0)
