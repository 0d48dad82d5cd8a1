"""One tracked frame, device resident (what a tracker calls per frame): extract -> frame set (undistort + grid) ->
SearchByProjection against the previous frame's points.  Wall clock per stage, one 640x480 frame, 1000 features."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import oracle as O
from orb_slam_b200 import Camera, FrameSet, KP_DTYPE, Matcher, ORBextractor, ORBmatcher
from orb_slam_b200.matcher import ProjectedPoints
H, W, NF = 480, 640, 1000
img = torch.from_numpy(O.synth(H, W, 1)).cuda()
ex = ORBextractor(NF, max_width=W, max_height=H, max_batch=1)
m = Matcher(0)
cam = Camera(517.3, 516.5, 318.6, 255.3, 0.2624, -0.9531, -0.0054, 0.0026, 1.1633, W, H)
d_k = torch.zeros((1, NF, 28), dtype=torch.uint8, device="cuda"); d_d = torch.zeros((1, NF, 32), dtype=torch.uint8, device="cuda")
d_n = torch.zeros(1, dtype=torch.int32, device="cuda")
st = torch.cuda.current_stream().cuda_stream
def extract():
    ex.extract_batch_device(img.data_ptr(), 1, W, H, W, H * W, d_k.data_ptr(), d_d.data_ptr(), d_n.data_ptr(), st)
    torch.cuda.synchronize()
sets = []
def glue():
    for s in sets: s.close()
    sets.clear(); sets.append(FrameSet.from_extractor(m, d_k.data_ptr(), d_d.data_ptr(), d_n.data_ptr(), 1, NF, cam, st))
def t(fn, reps=300):
    for _ in range(20): fn()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    return (time.perf_counter() - t0) / reps * 1e6
a = t(extract); b = t(glue)
fs = sets[0]
x, y, ang, octv, desc = fs.keypoints(0)[:5]
sf = np.cumprod(np.concatenate([[1.0], np.full(7, np.float32(1.2))]).astype(np.float32)).astype(np.float32)
pts = ProjectedPoints.for_last_frame(x + 3, y - 2, octv, desc, ang, sf, 15.0)
om = ORBmatcher(0.9, True, matcher=m)
c = t(lambda: om.SearchByProjection(fs, [0], [pts], "best", 100))
from oracle import bow_oracle as B
from orb_slam_b200 import ORBVocabulary
V = B.synth_vocabulary_fast(k=10, L=6, seed=3)
voc = ORBVocabulary.from_arrays(m, **{k: V[k] for k in ("k", "L", "scoring", "weighting", "parent", "desc", "weight", "is_leaf")})
d = t(lambda: fs.compute_bow(voc, 4))
om5 = ORBmatcher(0.75, True, matcher=m)
fs.set_valid(0, np.ones(fs.frame_n(0), np.uint8))
e = t(lambda: om5.SearchByBoW(fs, fs, [0], [0], "kf_f"))
print("ComputeBoW (k=10, L=6 vocabulary, 1 frame) %.1f us | SearchByBoW against itself %.1f us" % (d, e))
print("extract (device in/out) %.1f us | frame set from extractor %.1f us | SearchByProjection 1000 points %.1f us | step %.1f us" % (a, b, c, a + b + c))
