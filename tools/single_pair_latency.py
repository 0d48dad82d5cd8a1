"""Latency of the matcher calls the tracker makes once per frame, for ONE frame pair (1000 keypoints each)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from oracle import oracle as O
import matchdata as MD
from orb_slam_b200 import Matcher, ORBmatcher
from orb_slam_b200.matcher import ProjectedPoints
m = Matcher(0)
f1, f2, (dx, dy) = MD.make_pair(O, 3)
F1, F2 = MD.to_matchframe(f1), MD.to_matchframe(f2)
def t(fn, reps=200):
    for _ in range(10): fn()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    return (time.perf_counter() - t0) / reps * 1e6
sets = []
def up():
    for s in sets: s.close()
    sets.clear(); sets.append(m.upload([F1, F2]))
print("upload 2 frames (frameset create + grid): %.1f us" % t(up))
fs = m.upload([F1, F2])
om = ORBmatcher(0.9, True, matcher=m)
prev = [np.stack([f1.x, f1.y], 1).astype(np.float32)]
print("SearchForInitialization, 1 pair: %.1f us (kernels %.1f us)" % (t(lambda: om.SearchForInitialization(fs, fs, [0], [1], [prev[0].copy()], 100)), m.last_kernel_ms() * 1e3))
sf = np.cumprod(np.concatenate([[1.0], np.full(7, np.float32(1.2))]).astype(np.float32)).astype(np.float32)
pts = ProjectedPoints.for_last_frame(f1.x - dx, f1.y - dy, f1.octave, f1.desc, f1.angle, sf, 15.0)   # true motion of the pair
om2 = ORBmatcher(0.9, True, matcher=m)
print("SearchByProjection (last frame), 1 problem: %.1f us (kernels %.1f us)" % (t(lambda: om2.SearchByProjection(fs, [1], [pts], "best", 100)), m.last_kernel_ms() * 1e3))
om3 = ORBmatcher(0.75, True, matcher=m)
fs.set_valid(0, np.ones(F1.n, np.uint8))
print("SearchByBoW kf_f, 1 pair: %.1f us (kernels %.1f us)" % (t(lambda: om3.SearchByBoW(fs, fs, [0], [1], "kf_f")), m.last_kernel_ms() * 1e3))
q = f1.desc; db = f2.desc
print("knn2 1000 x 1000: %.1f us (kernels %.1f us)" % (t(lambda: m.knn2(q, db)), m.last_kernel_ms() * 1e3))
# SearchForTriangulation, one keyframe pair (LocalMapping::CreateNewMapPoints, one neighbour)
fv1 = MD.with_valid(f1, MD.random_valid(f1.n, 0.3, 60)); fv2 = MD.with_valid(f2, MD.random_valid(f2.n, 0.3, 70))
fst = m.upload([MD.to_matchframe(fv1), MD.to_matchframe(fv2)])
tvec = np.array([dx, dy, 0.0]) + 1e-3
F12 = np.array([[0, -tvec[2], tvec[1]], [tvec[2], 0, -tvec[0]], [-tvec[1], tvec[0], 0]], np.float32) * np.float32(0.05)
om4 = ORBmatcher(0.6, False, matcher=m)
print("SearchForTriangulation, 1 pair: %.1f us (kernels %.1f us)" % (t(lambda: om4.SearchForTriangulation(fst, fst, [0], [1], F12[None], (sf * sf).astype(np.float32))), m.last_kernel_ms() * 1e3))
