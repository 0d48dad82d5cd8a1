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


def _point_cloud(rng, rf, du, dv, cam, sf, nlevels=6):
    """World points (camera frame == world frame) for the MapPoints of frame rf so that they project near
    (x + du, y + dv); some behind the camera, outside the image, out of the scale range or seen from behind."""
    n = rf.n
    fx, fy, cx, cy = (np.float32(c) for c in cam[:4])
    z = rng.uniform(1.0, 8.0, n).astype(np.float32)
    z[5:10] = -2.0
    tu = rf.xu + du + rng.normal(0, 2, n); tv = rf.yu + dv + rng.normal(0, 2, n)
    tu[:5] = -400.0
    pos = np.stack([(tu - cx) / fx * z, (tv - cy) / fy * z, z], 1).astype(np.float32)
    dist = _norm3(pos)
    lvl_true = np.clip(rf.kps["octave"] + rng.integers(-1, 2, n), 0, nlevels - 1)
    min_dist = (dist / (sf[lvl_true] * rng.uniform(0.85, 0.999, n).astype(np.float32))).astype(np.float32)
    max_dist = (min_dist * np.float32(1.2 ** nlevels)).astype(np.float32)
    min_dist[10:14] = dist[10:14] * np.float32(1.5)
    max_dist[14:18] = dist[14:18] * np.float32(0.5)
    normal = (pos / dist[:, None]).astype(np.float32)
    flip = rng.random(n) < 0.1
    normal[flip] = -normal[flip]
    return pos, dist, min_dist, max_dist, normal


def _project_kf(pos, dist, min_dist, max_dist, normal, cam, bounds, sf, check_normal):
    """Caller-side float expressions shared by Fuse (src/ORBmatcher.cc:1043-1083, :1171-1212) and SearchBySim3
    (:1325-1357, :1410-1442) at identity pose: (u, v, predicted level, passes-all-tests)."""
    from orb_slam_b200.matcher import predict_scale_level
    fx, fy, cx, cy = (np.float32(c) for c in cam[:4])
    b = bounds
    invz = (np.float32(1) / pos[:, 2]).astype(np.float32)
    x = (pos[:, 0] * invz).astype(np.float32); y = (pos[:, 1] * invz).astype(np.float32)
    u = fx * x + cx; v = fy * y + cy
    in_img = (u >= np.float32(b[0])) & (u < np.float32(b[1])) & (v >= np.float32(b[2])) & (v < np.float32(b[3]))
    ok = (~(pos[:, 2] < 0)) & in_img & ~((dist < min_dist) | (dist > max_dist))
    if check_normal:
        dot = (pos.astype(np.float64) * normal.astype(np.float64)).sum(1)
        ok &= ~(dot < 0.5 * dist.astype(np.float64))
    return u, v, predict_scale_level(dist, min_dist, sf), ok


def fuse_cases(oracle, RF, seed, cam, nfeatures=400):
    """ORBmatcher::Fuse (both overloads), SearchBySim3 and SearchByProjection(F1, F2, window) run by the COMPILED
    REFERENCE on its own objects.  Yields dicts: `kind` in {fuse, fuse_sim3, sim3, proj_ff}, the flat-array inputs of the
    oracle / CUDA form and everything the reference's call leaves observable."""
    from orb_slam_b200.matcher import ProjectedPoints
    w = RF.World(None, nfeatures, 1.2, 6, 1, 20)
    ims = images(oracle, seed)
    f1 = w.frame(ims[0], K_of(cam), cam[4:9], reset_static=True)
    f2 = w.frame(ims[1], K_of(cam), cam[4:9])
    sf = np.cumprod(np.concatenate([[1.0], np.full(5, np.float32(1.2))]).astype(np.float32)).astype(np.float32)
    rng = np.random.default_rng(seed)
    v1, v2 = valid_flags(f1.n, seed + 2), valid_flags(f2.n, seed + 1)
    g1 = _point_cloud(rng, f1, -11, -6, cam, sf)      # f1's points, seen in f2
    g2 = _point_cloud(rng, f2, +11, +6, cam, sf)      # f2's points, seen in f1

    def fresh():
        """New MapPoint objects on both frames (the calls below mutate them)."""
        f1.set_valid(v1); f1.set_point_geometry(g1[0], g1[2], g1[3], g1[4]); f1.points_take_descriptors()
        f2.set_valid(v2); f2.set_point_geometry(g2[0], g2[2], g2[3], g2[4]); f2.points_take_descriptors()

    def frame_of(rf, v):
        t = match_frame(rf, v, None)
        t.valid = (np.asarray(v) != 0).astype(np.uint8)
        return t

    sel = np.flatnonzero(v1 != 0).astype(np.int32)     # every MapPoint of f1, good and bad
    rng.shuffle(sel)
    # ---- Fuse(pKF, vpMapPoints, th) :1016-1134 ------------------------------------------------------------------------
    u, v, level, ok = _project_kf(*g1, cam, f2.bounds, sf, True)
    for th in (2.5, 6.0):
        fresh()
        n, added, went_bad, repl = RF.fuse(f2, f1, sel, th)
        pts = ProjectedPoints.for_fuse(u[sel], v[sel], level[sel], f1.desc[sel], sf, th, valid=(v1[sel] == 1) & ok[sel])
        yield dict(kind="fuse", name=f"fuse:{th}", target=frame_of(f2, v2), tvalid=v2.copy(), points=pts, th_dist=50,
                   ref_n=n, added=added, went_bad=went_bad, repl=repl)
    # ---- Fuse(pKF, Scw, vpPoints, th) :1136-1265 ------------------------------------------------------------------------
    for th in (4.0, 8.0):
        fresh()
        n, added, tbad = RF.fuse_sim3(f2, f1, sel, th)
        pts = ProjectedPoints.for_fuse(u[sel], v[sel], level[sel], f1.desc[sel], sf, th, valid=(v1[sel] == 1) & ok[sel])
        yield dict(kind="fuse_sim3", name=f"fuse_sim3:{th}", target=frame_of(f2, v2), tvalid=v2.copy(), points=pts,
                   th_dist=50, ref_n=n, added=added, target_went_bad=tbad)
    # ---- SearchBySim3 :1267-1505 ----------------------------------------------------------------------------------------
    u12, v12, l12, ok12 = _project_kf(*g1, cam, f2.bounds, sf, False)
    u21, v21, l21, ok21 = _project_kf(*g2, cam, f1.bounds, sf, False)
    for th in (7.5, 3.0):
        fresh()
        pre = np.full(f1.n, -1, np.int32)              # vpMatches12 on entry: some pairs already matched
        c1 = rng.choice(np.flatnonzero(v1 == 1), 20, replace=False)
        pre[c1] = rng.choice(np.flatnonzero(v2 != 0), 20, replace=False)
        n, out12 = RF.search_by_sim3(f1, f2, pre, th)
        already2 = np.zeros(f2.n, bool); already2[pre[pre >= 0]] = True
        p12 = ProjectedPoints.for_fuse(u12, v12, l12, f1.desc, sf, th, valid=(v1 == 1) & (pre < 0) & ok12)
        p21 = ProjectedPoints.for_fuse(u21, v21, l21, f2.desc, sf, th, valid=(v2 == 1) & ~already2 & ok21)
        yield dict(kind="sim3", name=f"sim3:{th}", f1=frame_of(f1, v1), f2=frame_of(f2, v2), p12=p12, p21=p21, pre=pre,
                   th_dist=100, ref_n=n, ref_m=out12)
    # ---- SearchByProjection(F1, F2, window, matches2) :519-594 ----------------------------------------------------------
    fx, fy, cx, cy = (np.float32(c) for c in cam[:4])
    pos = g1[0]
    invz = (1.0 / pos[:, 2].astype(np.float64)).astype(np.float32)          # const float invzc2 = 1.0/z
    uu = (fx * pos[:, 0]) * invz + cx; vv = (fy * pos[:, 1]) * invz + cy     # F2.fx*xc2*invzc2+F2.cx, left to right
    for window, ratio in ((15, 0.8), (50, 0.9)):
        fresh()
        n, m21 = RF.search_by_projection_ff(f1, f2, pos, window, ratio, True)
        a = match_frame(f1, v1, None)                   # valid = good MapPoint (none of them is in F2 yet)
        yield dict(kind="proj_ff", name=f"proj_ff:{window}", f1=a, f2=frame_of(f2, v2),
                   uv=np.stack([uu, vv], 1).astype(np.float32), window=window, ratio=ratio, ref_n=n, ref_m=m21)


def fuse_expectation(case, best):
    """What the reference's Fuse leaves observable, predicted from the per-point best keypoints `best` (oracle or CUDA)
    and the state of the target keypoints (`tvalid`: 0 no MapPoint, 1 good, 2 bad) by replaying ONLY the pointer
    bookkeeping of src/ORBmatcher.cc:1116-1131 / :1245-1260 and MapPoint::Replace (src/MapPoint.cc:124-158)."""
    tv = case["tvalid"]
    n = len(best)
    if case["kind"] == "fuse":
        added = np.full(n, -1, np.int32); bad = np.zeros(n, np.uint8); repl = np.full(n, -1, np.int32)
        holder = {}                     # keypoint -> 'own' (the keyframe's good MapPoint) or a point index
        seen_src = set()                # MapPoints in the keyframe that already observe the source keyframe
        for k in range(n):
            t = int(best[k])
            if t < 0:
                continue
            h = holder.get(t, "own" if tv[t] == 1 else ("badown" if tv[t] == 2 else None))
            if h is None:               # free keypoint: AddObservation + AddMapPoint
                added[k] = t; holder[t] = k; seen_src.add(("p", k))
            elif h == "badown":         # a bad MapPoint sits there: counted, nothing done
                pass
            else:                       # pMP->Replace(pMPinKF): the point dies, its source observation moves over
                bad[k] = 1
                key = ("own", t) if h == "own" else ("p", h)
                if key not in seen_src:
                    seen_src.add(key)
                    if h == "own":
                        repl[k] = t     # observable through the keyframe's own frame index
        return dict(added=added, went_bad=bad, repl=repl)
    added = np.full(n, -1, np.int32); tbad = np.zeros(len(tv), np.uint8)
    last = {}
    for k in range(n):
        t = int(best[k])
        if t >= 0 and tv[t] != 2:
            last[t] = k                 # pMPinKF->Replace(pMP): the newest point takes over the observation
            if tv[t] == 1:
                tbad[t] = 1
    for t, k in last.items():
        added[k] = t
    return dict(added=added, target_went_bad=tbad)
