import os, sys
import numpy as np
sys.path.insert(0, os.getcwd())
from orb_slam_b200 import ORBextractor
rng = np.random.default_rng(5)
noise = rng.integers(0, 256, (480, 640), dtype=np.uint8)
flat = np.full((480, 640), 77, np.uint8)
half = flat.copy(); half[:, 320:] = noise[:, 320:]
def mk(mode):
    os.environ["ORBX_FAST"] = mode
    return ORBextractor(max_batch=1)
A, B = mk("legacy"), mk("tma")
for name, img in (("noise", noise), ("flat", flat), ("half", half)):
    A(img); B(img)
    nd = 0
    for l in range(8):
        info = A.debug_level_info(l)
        ta, _ = A.debug_level_cells(0, l); tb, _ = B.debug_level_cells(0, l)
        for c in range(info["cols"] * info["rows"]):
            xa, ya, ra = A.debug_cell_candidates(0, l, c); xb, yb, rb = B.debug_cell_candidates(0, l, c)
            sa = sorted(zip(ya.tolist(), xa.tolist(), ra.tolist())); sb = sorted(zip(yb.tolist(), xb.tolist(), rb.tolist()))
            if ta[c] != tb[c] or sa != sb:
                nd += 1
                if nd <= 6:
                    only_a = sorted(set(sa) - set(sb))[:5]; only_b = sorted(set(sb) - set(sa))[:5]
                    print(name, "level", l, "cell", c, "n", ta[c], tb[c], "onlyA", only_a, "onlyB", only_b, "dupB", len(sb) - len(set(sb)))
    print(name, "cells differing:", nd)
