"""Shared test inputs for the matcher: frame pairs extracted by the CPU oracle from shifted synthetic images,
synthetic DBoW2-like feature vectors, projected points.  Test infrastructure (imports oracle/)."""
from __future__ import annotations

from types import SimpleNamespace

import numpy as np


def frame_from(kps, desc, w, h, valid=None, featvec=None):
    return SimpleNamespace(x=kps["x"].copy(), y=kps["y"].copy(), angle=kps["angle"].copy(),
                           octave=kps["octave"].copy(), desc=desc.copy(), valid=valid, min_x=0.0, min_y=0.0,
                           max_x=float(w), max_y=float(h), featvec=featvec, n=len(kps))


def synth_featvec(desc, nnodes=40, seed=0):
    """Deterministic stand-in for ORBVocabulary::transform(..., levelsup=4): bucket = hash of the leading descriptor
    bits into <= nnodes sparse node ids (the vocabulary file is absent)."""
    key = (desc[:, 0].astype(np.int64) * 131 + (desc[:, 1] >> 4) + seed) % nnodes
    node_of = key * 7 + 3  # sparse, non-contiguous ids
    ids = np.unique(node_of)
    start = [0]
    items = []
    for i in ids:
        it = np.nonzero(node_of == i)[0]
        items.extend(it.tolist())
        start.append(len(items))
    return ids.astype(np.int32), np.asarray(start, np.int32), np.asarray(items, np.int32)


_cache = {}


def make_pair(oracle, seed, w=640, h=480, nfeat=1000, shift=None, with_fv=True):
    """Two views of one synthetic scene: view 2 is translated by `shift` px with fresh +-2 noise."""
    key = (seed, w, h, nfeat, shift, with_fv)
    if key in _cache:
        return _cache[key]
    rng = np.random.default_rng(1000 + seed)
    if shift is None:
        shift = (int(rng.integers(-30, 31)), int(rng.integers(-30, 31)))
    m = 48
    big = oracle.synth(h + 2 * m, w + 2 * m, seed)
    im1 = np.ascontiguousarray(big[m:m + h, m:m + w])
    dx, dy = shift
    im2 = big[m + dy:m + dy + h, m + dx:m + dx + w].astype(np.int16) + rng.integers(-2, 3, (h, w))
    im2 = np.ascontiguousarray(np.clip(im2, 0, 255).astype(np.uint8))
    E = oracle.OracleExtractor(nfeat, 1.2, 8, 1, 20)
    k1, d1 = E(im1)
    k2, d2 = E(im2)
    f1 = frame_from(k1, d1, w, h, featvec=synth_featvec(d1) if with_fv else None)
    f2 = frame_from(k2, d2, w, h, featvec=synth_featvec(d2) if with_fv else None)
    _cache[key] = (f1, f2, shift)
    return _cache[key]


def with_valid(f, valid):
    g = SimpleNamespace(**vars(f))
    g.valid = np.ascontiguousarray(valid, np.uint8)
    return g


def random_valid(n, frac, seed):
    return (np.random.default_rng(seed).random(n) < frac).astype(np.uint8)


def to_matchframe(f):
    from orb_slam_b200 import MatchFrame
    return MatchFrame(f.x, f.y, f.angle, f.octave, f.desc, f.valid, f.min_x, f.min_y, f.max_x, f.max_y, f.featvec)


def frustum_scene(seed, n=5000, w=640, h=480, nlevels=8):
    """A random camera pose and map points spread so that every exit of Frame::isInFrustum is taken: behind the
    camera, outside the image, too near / too far, grazing view, and accepted points over all predicted levels.
    Returns (Rcw, tcw, intrinsics, bounds, scale_factors, pos, normal, min_dist, max_dist), all float32."""
    rng = np.random.default_rng(seed)
    q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
    if np.linalg.det(q) < 0:
        q[:, 0] = -q[:, 0]
    Rcw = q.astype(np.float32)
    tcw = rng.normal(size=3).astype(np.float32) * 2
    fx, fy, cx, cy = np.float32(525.0), np.float32(523.5), np.float32(w / 2 - 3.5), np.float32(h / 2 + 1.25)
    sf = np.cumprod(np.concatenate([[1.0], np.full(nlevels - 1, np.float32(1.2))]).astype(np.float32)).astype(np.float32)
    # points in camera coordinates, then moved to the world
    z = rng.uniform(-2, 30, n)
    x = rng.uniform(-0.9, 0.9, n) * np.abs(z) * (w / 2) / 525.0
    y = rng.uniform(-0.9, 0.9, n) * np.abs(z) * (h / 2) / 523.5
    pc = np.stack([x, y, z], 1)
    pos = ((pc - tcw.astype(np.float64)) @ Rcw.astype(np.float64)).astype(np.float32)   # R^T (pc - t)
    Ow = -(Rcw.astype(np.float64).T @ tcw.astype(np.float64))
    po = pos.astype(np.float64) - Ow
    dist = np.linalg.norm(po, axis=1)
    nrm = po / np.maximum(dist[:, None], 1e-9) + rng.normal(size=(n, 3)) * 0.8
    nrm = (nrm / np.linalg.norm(nrm, axis=1, keepdims=True)).astype(np.float32)
    mn = (dist / rng.uniform(0.7, 6.0, n)).astype(np.float32)
    mx = (mn * rng.uniform(1.0, 5.0, n)).astype(np.float32)
    return Rcw, tcw, (fx, fy, cx, cy), (0, w, 0, h), sf, pos, nrm, mn, mx
