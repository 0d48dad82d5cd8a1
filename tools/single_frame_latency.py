"""Config 1 (one 640x480 frame per call) latency breakdown: wall clock of ORBextractor.__call__ (pageable numpy),
of the pinned-buffer C-ABI call, and the serial per-stage kernel times of one frame."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import oracle as O
from orb_slam_b200 import ORBextractor, KP_DTYPE
H, W, NF = 480, 640, 1000
img = O.synth(H, W, 1)
ex = ORBextractor(NF, max_width=W, max_height=H, max_batch=1)
for _ in range(20): ex(img)
t = time.perf_counter()
for _ in range(200): ex(img)
print("ORBextractor.__call__ (pageable): %.1f us" % ((time.perf_counter() - t) / 200 * 1e6))
pin = torch.from_numpy(img).pin_memory()
k = torch.zeros((NF, 28), dtype=torch.uint8).pin_memory(); d = torch.zeros((NF, 32), dtype=torch.uint8).pin_memory()
n = torch.zeros(1, dtype=torch.int32).pin_memory()
for _ in range(20): ex.extract_batch_ptr(pin.data_ptr(), 1, W, H, W, W * H, k.data_ptr(), d.data_ptr(), n.data_ptr())
t = time.perf_counter()
for _ in range(500): ex.extract_batch_ptr(pin.data_ptr(), 1, W, H, W, W * H, k.data_ptr(), d.data_ptr(), n.data_ptr())
print("orbx_extract_batch, 1 frame, pinned: %.1f us" % ((time.perf_counter() - t) / 500 * 1e6))
ex.profile_enable(True)
for _ in range(50): ex.extract_batch_ptr(pin.data_ptr(), 1, W, H, W, W * H, k.data_ptr(), d.data_ptr(), n.data_ptr())
ex.profile_read()
for _ in range(200): ex.extract_batch_ptr(pin.data_ptr(), 1, W, H, W, W * H, k.data_ptr(), d.data_ptr(), n.data_ptr())
pr = ex.profile_read()
print("serial stage us per frame:", {s: round(v[0] / 200 * 1e3, 1) for s, v in pr.items()}, "sum %.1f" % (sum(v[0] for v in pr.values()) / 200 * 1e3))
