"""Random-shape sweep of the tensor-core 2-NN (knn2_tc_kernel, forced with ORBM_KNN_TC=1) against the POPC kernels
(ORBM_KNN_TC=0) on the same device buffers, and against the CPU oracle for the smaller cases.  Keys are integers: every
comparison is bit-exact.   python tools/fuzz_knn2_tc.py [cases] [seed]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def handle(tc):
    from orb_slam_b200 import Matcher
    os.environ["ORBM_KNN_TC"] = "1" if tc else "0"
    try:
        return Matcher(0)
    finally:
        del os.environ["ORBM_KNN_TC"]


def main(cases, seed):
    import torch
    from oracle import match_oracle as MO
    MO.lib()
    rng = np.random.default_rng(seed)
    tc, pop = handle(True), handle(False)
    st = torch.cuda.current_stream().cuda_stream
    bad = checked_cpu = 0
    pairs = 0.0
    for c in range(cases):
        nq = int(np.exp(rng.uniform(0, np.log(4000))))
        ndb = int(np.exp(rng.uniform(0, np.log(300000))))
        kind = rng.integers(0, 4)
        db = rng.integers(0, 256, (ndb, 32), dtype=np.uint8)
        if kind == 1:                                    # few distinct descriptors: ties everywhere
            db = db[rng.integers(0, min(ndb, 7), ndb)]
        q = db[rng.integers(0, ndb, nq)] ^ np.packbits(rng.random((nq, 256)) < rng.choice([0.0, 0.03, 0.1, 0.5]), axis=1)
        if kind == 2:                                    # far side: most dot products negative
            q = ~q
        if kind == 3 and nq > 4:                         # a mix of near and far queries inside one warp
            q[::2] = ~q[::2]
        base = int(rng.integers(0, (1 << 23) - ndb))
        d_q, d_db = torch.from_numpy(q).cuda(), torch.from_numpy(db).cuda()
        a = torch.zeros((nq, 2), dtype=torch.int32, device="cuda"); b = torch.zeros_like(a)
        tc.knn2_device(d_q.data_ptr(), nq, d_db.data_ptr(), ndb, base, a.data_ptr(), st)
        pop.knn2_device(d_q.data_ptr(), nq, d_db.data_ptr(), ndb, base, b.data_ptr(), st)
        torch.cuda.synchronize()
        ok = bool(torch.equal(a, b))
        if ok and nq * ndb <= 3e7:
            ok = np.array_equal(a.cpu().numpy().view(np.uint32), MO.knn2(q, db, idx_base=base, threads=8)); checked_cpu += 1
        pairs += float(nq) * ndb
        if not ok:
            bad += 1
            print(f"MISMATCH case {c}: nq {nq} ndb {ndb} kind {kind} base {base}", flush=True)
    tc.synchronize(); pop.synchronize()
    print(f"{cases} cases ({checked_cpu} also against the CPU oracle), {pairs:.3g} pairs, mismatches: {bad}")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main(int(sys.argv[1]) if len(sys.argv) > 1 else 300, int(sys.argv[2]) if len(sys.argv) > 2 else 1))
