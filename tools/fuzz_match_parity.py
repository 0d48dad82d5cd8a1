"""Random-configuration parity sweep of the CUDA matcher against the CPU oracle (test infrastructure, like tests/).

  python tools/fuzz_match_parity.py [--rounds 40] [--seed 0]

Every round builds 1-6 frame pairs of random size / feature count / inter-frame shift with the oracle extractor, then
runs SearchForInitialization (two consecutive windows, updated vbPrevMatched carried over), WindowSearch (random
MapPoint masks and level windows) and the frame-to-frame SearchByProjection through the C ABI in ONE batched call each
and compares match counts and match rows with `oracle/match_oracle.cc` problem by problem.  Exit code 1 on a mismatch.
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import matchdata as MD                                # noqa: E402
from oracle import match_oracle as M                  # noqa: E402
from oracle import oracle as O                        # noqa: E402
from orb_slam_b200 import Matcher, ORBmatcher         # noqa: E402


def rows_equal(row, ref):
    return np.array_equal(row[: len(ref)], ref) and np.all(row[len(ref):] == -1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rounds", type=int, default=40)
    ap.add_argument("--seed", type=int, default=0)
    a = ap.parse_args()
    rng = np.random.default_rng(a.seed)
    O.lib(); M.lib()
    matcher = Matcher(0)
    bad = problems = matches = 0
    for r in range(a.rounds):
        npairs = int(rng.integers(1, 7))
        pairs = []
        while len(pairs) < npairs:
            w, h = int(rng.integers(200, 1300)), int(rng.integers(160, 700))
            nf = int(rng.choice([300, 700, 1000, 2000]))
            shift = (int(rng.integers(-40, 41)), int(rng.integers(-30, 31)))
            try:
                pairs.append(MD.make_pair(O, int(rng.integers(0, 1 << 20)), w, h, nf, shift=shift, with_fv=False))
            except RuntimeError:                      # a size / feature count the reference's cell grid cannot take
                continue
        idx = list(range(npairs))
        ratio, ori = float(rng.choice([0.6, 0.75, 0.9])), bool(rng.integers(0, 2))
        tag = f"round {r} ({npairs} pairs, ratio {ratio}, ori {ori})"
        om = ORBmatcher(ratio, ori, matcher=matcher)
        # --- SearchForInitialization, twice (the second call starts from the updated vbPrevMatched)
        s1 = om.upload([MD.to_matchframe(p[0]) for p in pairs]); s2 = om.upload([MD.to_matchframe(p[1]) for p in pairs])
        prev = [np.stack([p[0].x, p[0].y], 1).astype(np.float32) for p in pairs]
        oprev = [x.copy() for x in prev]
        for win in (int(rng.integers(30, 150)), int(rng.integers(5, 40))):
            nm, rows = om.SearchForInitialization(s1, s2, idx, idx, prev, win)
            for k, (f1, f2, _) in enumerate(pairs):
                on, orow = M.search_for_initialization(f1, f2, oprev[k], win, ratio, ori)
                problems += 1; matches += on
                if nm[k] != on or not rows_equal(rows[k], orow) or not np.array_equal(prev[k], oprev[k]):
                    bad += 1
                    print(f"FAIL {tag}: SearchForInitialization window {win} pair {k}: {nm[k]} vs {on}", flush=True)
        # --- WindowSearch with MapPoint masks and a level window
        fr1 = [MD.with_valid(p[0], MD.random_valid(p[0].n, float(rng.uniform(0.1, 1.0)), 10 * r + k)) for k, p in enumerate(pairs)]
        fr2 = [MD.with_valid(p[1], MD.random_valid(p[1].n, float(rng.uniform(0.0, 0.6)), 10 * r + k + 5)) for k, p in enumerate(pairs)]
        t1 = matcher.upload([MD.to_matchframe(f) for f in fr1]); t2 = matcher.upload([MD.to_matchframe(f) for f in fr2])
        win = int(rng.integers(10, 250))
        lo = int(rng.integers(0, 4)); hi = int(rng.choice([lo, lo + 2, 2**31 - 1]))
        nm, rows = om.WindowSearch(t1, t2, idx, idx, win, lo, hi)
        for k in idx:
            on, orow = M.window_search(fr1[k], fr2[k], win, ratio, ori, lo, hi)
            problems += 1; matches += on
            if nm[k] != on or not rows_equal(rows[k], orow):
                bad += 1
                print(f"FAIL {tag}: WindowSearch window {win} levels [{lo},{hi}] pair {k}: {nm[k]} vs {on}", flush=True)
        # --- frame-to-frame SearchByProjection around predicted positions
        win = int(rng.integers(5, 60))
        uv = [np.stack([f.x - np.float32(p[2][0]) + np.float32(rng.uniform(-2, 2)),
                        f.y - np.float32(p[2][1]) + np.float32(rng.uniform(-2, 2))], 1).astype(np.float32)
              for f, p in zip(fr1, pairs)]
        nm, rows = om.SearchByProjectionFrames(t1, t2, idx, idx, uv, win)
        for k in idx:
            on, orow = M.search_by_projection_ff(fr1[k], fr2[k], uv[k], win, ratio)
            problems += 1; matches += on
            if nm[k] != on or not rows_equal(rows[k], orow):
                bad += 1
                print(f"FAIL {tag}: SearchByProjectionFrames window {win} pair {k}: {nm[k]} vs {on}", flush=True)
    print(f"{a.rounds} rounds, {problems} problems, {matches} matches compared, {bad} failures")
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
