import faulthandler, sys, os, time
faulthandler.dump_traceback_later(50, exit=True)
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from oracle import match_oracle as M
M.lib()
from orb_slam_b200 import Matcher
os.environ["ORBM_KNN_TC"] = "1"; tc = Matcher(0)
os.environ["ORBM_KNN_TC"] = "0"; pop = Matcher(0)
rng = np.random.default_rng(2024)
shapes = [(1, 1), (1, 255), (127, 256), (128, 257), (129, 511), (255, 512), (256, 513), (257, 1000), (385, 3000),
          (5000, 300), (3000, 40), (60, 70000), (2, 150000), (1000, 20011)]
for nq, ndb in shapes:
    db = rng.integers(0, 256, (ndb, 32), dtype=np.uint8)
    q = db[rng.integers(0, ndb, nq)].copy()
    q ^= np.packbits(rng.random((nq, 256)) < 0.07, axis=1)
    q[::5] = db[rng.integers(0, ndb, len(q[::5]))]
    print("shape", nq, ndb, flush=True)
    ref = M.knn2(q, db, idx_base=37, threads=8); print("  oracle done", flush=True)
    got = tc.knn2(q, db, idx_base=37); print("  tc done", np.array_equal(got, ref), flush=True)
    g2 = pop.knn2(q, db, idx_base=37); print("  pop done", np.array_equal(g2, ref), flush=True)
db = np.tile(rng.integers(0, 256, (1, 32), dtype=np.uint8), (3000, 1))
q = rng.integers(0, 256, (300, 32), dtype=np.uint8)
print("identical", flush=True)
got = tc.knn2(q, db); print("  tc done", np.array_equal(got, M.knn2(q, db, threads=8)), flush=True)
q = rng.integers(0, 256, (500, 32), dtype=np.uint8)
far = (~q[rng.integers(0, 500, 20000)]).copy()
far ^= np.packbits(rng.random((20000, 256)) < 0.05, axis=1)
far[123] = ~q[7]
print("far", flush=True)
ref = M.knn2(q, far, threads=8)
got = tc.knn2(q, far); print("  tc done", np.array_equal(got, ref), (ref[:, 0] >> 23).min(), flush=True)
mixed = far.copy(); mixed[15000:15250] = q[:250]
print("mixed", np.array_equal(tc.knn2(q, mixed), M.knn2(q, mixed, threads=8)), flush=True)
