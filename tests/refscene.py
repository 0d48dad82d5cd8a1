"""Scenes for the compiled-reference pins (test infrastructure): a pair of shifted views, built into the reference's own
Frame objects (oracle/ref_frontend.py) and, from what those hold, into the flat MatchFrame records the oracle and the
CUDA matcher take."""
from __future__ import annotations

import numpy as np

CAM_SHIPPED = (268.9633, 269.9858, 157.6087, 114.6369, -0.4157, 0.2624, 0.0, 0.0, -0.1178)   # Data/Settings.yaml
CAM_PINHOLE = (300.0, 300.0, 160.0, 120.0, 0.0, 0.0, 0.0, 0.0, 0.0)                           # k1 == 0: no undistortion


def images(oracle, seed, h=240, w=320, shift=(11, 6)):
    big = oracle.synth(h + 64, w + 64, seed)
    a = np.ascontiguousarray(big[32:32 + h, 32:32 + w])
    b = np.ascontiguousarray(big[32 + shift[1]:32 + shift[1] + h, 32 + shift[0]:32 + shift[0] + w])
    return a, b


def K_of(cam):
    return [cam[0], 0, cam[2], 0, cam[1], cam[3], 0, 0, 1]


def valid_flags(n, seed):
    rng = np.random.default_rng(seed)
    v = (rng.random(n) < 0.7).astype(np.uint8)
    v[rng.random(n) < 0.05] = 2          # a MapPoint whose isBad() is true
    return v


def match_frame(rf, valid, featvec):
    """MatchFrame from what the reference Frame holds (undistorted keypoints, bounds)."""
    from orb_slam_b200.matcher import MatchFrame
    b = rf.bounds
    return MatchFrame(rf.xu, rf.yu, rf.kps["angle"], rf.kps["octave"], rf.desc, (np.asarray(valid) == 1).astype(np.uint8),
                      float(b[0]), float(b[2]), float(b[1]), float(b[3]), featvec)


def _norm3(p):
    """cv::norm of a 3x1 float Mat: double accumulation of squares, sqrt, then the reference stores it in a float."""
    p = p.astype(np.float64)
    return np.sqrt(p[:, 0] * p[:, 0] + p[:, 1] * p[:, 1] + p[:, 2] * p[:, 2]).astype(np.float32)


def projection_cases(oracle, RF, seed, cam, nfeatures=400):
    """The SearchByProjection family run by the COMPILED REFERENCE on its own objects.  Yields one dict per
    call: the flat-array inputs of the oracle / CUDA form (`target` MatchFrame, `points` ProjectedPoints, variant,
    th_dist, ratio, ori) and what the reference returned (`ref_n`, `ref_m`).  The caller-side float expressions of each
    overload (projection, bounds, level prediction) are restated here in numpy next to the line they mirror."""
    from orb_slam_b200.matcher import ProjectedPoints, predict_scale_level
    w = RF.World(None, nfeatures, 1.2, 6, 1, 20)
    ims = images(oracle, seed)
    f1 = w.frame(ims[0], K_of(cam), cam[4:9], reset_static=True)        # source of the map points
    f2 = w.frame(ims[1], K_of(cam), cam[4:9])                           # current frame / loop keyframe
    sf = np.cumprod(np.concatenate([[1.0], np.full(5, np.float32(1.2))]).astype(np.float32)).astype(np.float32)
    rng = np.random.default_rng(seed)
    fx, fy, cx, cy = (np.float32(cam[0]), np.float32(cam[1]), np.float32(cam[2]), np.float32(cam[3]))
    b = f2.bounds
    n = f1.n
    v1, v2 = valid_flags(n, seed + 2), valid_flags(f2.n, seed + 1)

    def target(valid_any):
        t = match_frame(f2, v2, None)
        t.valid = np.asarray(valid_any, np.uint8)       # the searches skip a keypoint holding ANY MapPoint pointer
        return t

    # ---- local map (:49-125): one point per keypoint of f1, projected near where the scene moved to in f2 -------
    for th, ratio in ((1.0, 0.8), (5.0, 0.8), (3.0, 0.6)):
        f2.set_valid(v2)
        u = (f1.xu - 11 + rng.normal(0, 1.5, n)).astype(np.float32)
        v = (f1.yu - 6 + rng.normal(0, 1.5, n)).astype(np.float32)
        level = np.clip(f1.kps["octave"] + rng.integers(-1, 2, n), 0, 5).astype(np.int32)
        vcos = rng.choice(np.array([0.6, 0.9, 0.998, 0.9985, 1.0], np.float32), n)
        in_view = (rng.random(n) < 0.9).astype(np.uint8)
        bad = (rng.random(n) < 0.05).astype(np.uint8)
        rn, rm = RF.search_by_projection_local(f2, f1, u, v, level, vcos, in_view, bad, np.arange(n, dtype=np.int32),
                                               th, ratio)
        pts = ProjectedPoints.for_local_map(u, v, vcos, level, f1.desc, sf, th, valid=(in_view == 1) & (bad == 0))
        yield dict(name=f"local_map:{th}:{ratio}", target=target(v2 != 0), points=pts, variant="local_map",
                   th_dist=100, ratio=ratio, ori=True, ref_n=rn, ref_m=rm)

    # ---- world points (camera frame == world frame) for the three pose-taking overloads -------------------------
    z = rng.uniform(1.0, 8.0, n).astype(np.float32)
    z[5:10] = -2.0                                     # behind the camera (Sim3 overload: skipped, :325-326)
    tu = f1.xu - 11 + rng.normal(0, 2, n); tv = f1.yu - 6 + rng.normal(0, 2, n)
    tu[:5] = -400.0                                    # projected outside the image
    pos = np.stack([(tu - cx) / fx * z, (tv - cy) / fy * z, z], 1).astype(np.float32)
    dist = _norm3(pos)                                 # |x3Dw - Ow|, Ow = 0
    lvl_true = np.clip(f1.kps["octave"] + rng.integers(-1, 2, n), 0, 5)
    min_dist = (dist / (sf[lvl_true] * rng.uniform(0.85, 0.999, n).astype(np.float32))).astype(np.float32)
    max_dist = (min_dist * np.float32(1.2 ** 6)).astype(np.float32)
    min_dist[10:14] = dist[10:14] * np.float32(1.5)    # closer than minDistance (Sim3: skipped; reloc: level 0)
    max_dist[14:18] = dist[14:18] * np.float32(0.5)    # farther than maxDistance (Sim3: skipped)
    normal = (pos / dist[:, None]).astype(np.float32)
    flip = rng.random(n) < 0.1
    normal[flip] = -normal[flip]                       # seen from behind: PO.dot(Pn) < 0.5 dist (Sim3: skipped)
    f1.set_valid(v1)
    f1.set_point_geometry(pos, min_dist, max_dist, normal)
    f1.points_take_descriptors()
    level = predict_scale_level(dist, min_dist, sf)    # :1660-1667 / :355-360
    assert len(set(level.tolist())) >= 5

    invz = (1.0 / pos[:, 2].astype(np.float64)).astype(np.float32)          # const float invzc = 1.0/z
    uu = (fx * pos[:, 0]) * invz + cx                                        # :1535-1538 / :1652-1653, left to right
    vv = (fy * pos[:, 1]) * invz + cy
    inside = ~((uu < np.float32(b[0])) | (uu > np.float32(b[1])) | (vv < np.float32(b[2])) | (vv > np.float32(b[3])))

    # ---- last frame -> current frame (:1507-1620) ---------------------------------------------------------------
    for th, ratio, ori in ((15.0, 0.9, True), (7.0, 0.9, False)):
        f2.set_valid(v2)
        rn, rm = RF.search_by_projection_last(f2, f1, pos, th, ratio, ori)
        pts = ProjectedPoints.for_last_frame(uu, vv, f1.kps["octave"], f1.desc, f1.kps["angle"], sf, th,
                                             valid=(v1 != 0) & inside)       # :1527-1531: bad points are NOT skipped
        yield dict(name=f"last_frame:{th}:{ori}", target=target(v2 != 0), points=pts, variant="best", th_dist=100,
                   ratio=ratio, ori=ori, ref_n=rn, ref_m=rm)

    # ---- relocalisation (:1622-1746) ----------------------------------------------------------------------------
    already = (rng.random(n) < 0.15).astype(np.uint8)
    for th, orb_dist, ori in ((10.0, 100, True), (3.0, 64, True), (6.0, 100, False)):
        f2.set_valid(v2)
        rn, rm = RF.search_by_projection_reloc(f2, f1, already, th, orb_dist, 0.9, ori)
        pts = ProjectedPoints.for_relocalisation(uu, vv, level, f1.desc, f1.kps["angle"], sf, th,
                                                 valid=(v1 == 1) & (already == 0) & inside)
        yield dict(name=f"reloc:{th}:{orb_dist}:{ori}", target=target(v2 != 0), points=pts, variant="best",
                   th_dist=orb_dist, ratio=0.9, ori=ori, ref_n=rn, ref_m=rm)

    # ---- Sim3 / loop closing (:286-407) -------------------------------------------------------------------------
    x = (pos[:, 0] * (np.float32(1) / pos[:, 2])).astype(np.float32)         # :329-331: invz = 1/z in float
    y = (pos[:, 1] * (np.float32(1) / pos[:, 2])).astype(np.float32)
    us = fx * x + cx; vs = fy * y + cy
    in_img = (us >= np.float32(b[0])) & (us < np.float32(b[1])) & (vs >= np.float32(b[2])) & (vs < np.float32(b[3]))
    dot = (pos.astype(np.float64) * normal.astype(np.float64)).sum(1)        # Mat::dot: double accumulation
    ok = (~(pos[:, 2] < 0)) & in_img & ~((dist < min_dist) | (dist > max_dist)) & ~(dot < 0.5 * dist.astype(np.float64))
    sel = np.flatnonzero(v1 != 0).astype(np.int32)       # vpPoints: every MapPoint of f1 (good and bad)
    rng.shuffle(sel)
    for th in (10, 4):
        f2.set_valid(v2)
        pre = np.full(len(sel), -1, np.int32)            # points already matched to some keyframe keypoint
        k_pre = rng.choice(len(sel), 25, replace=False)
        pre[k_pre] = rng.choice(np.flatnonzero(v2 == 0), 25, replace=False)
        rn, rm = RF.search_by_projection_sim3(f2, f1, sel, pre, th, 0.9)
        taken = (v2 != 0)
        taken[pre[pre >= 0]] = True                      # vpMatched[idx] != NULL
        pts = ProjectedPoints.for_sim3(us[sel], vs[sel], level[sel], f1.desc[sel], sf, th,
                                       valid=(v1[sel] == 1) & (pre < 0) & ok[sel])
        yield dict(name=f"sim3:{th}", target=target(taken), points=pts, variant="best", th_dist=50, ratio=0.9,
                   ori=False, ref_n=rn, ref_m=rm)
