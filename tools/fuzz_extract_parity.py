"""Random-knob parity sweep of the CUDA extractor against the CPU oracle (test infrastructure, like tests/).

  python tools/fuzz_extract_parity.py [--cases 200] [--seed 0]

Each case draws image size, nfeatures, scaleFactor, nlevels, score type, FAST threshold, blur / retain mode and an
image kind (synthetic scene, uniform noise, low-contrast scene, blocks), runs 1-3 frames through `orbx_extract_batch`
and through `OracleExtractor`, and compares every keypoint field (angles as raw bits) and descriptor byte.  A
configuration the oracle calls degenerate (the reference would assert) must be refused by `orbx_create` or
`orbx_extract` too.  Prints one line per failure and a summary; exit code 1 on any mismatch.
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle as O                      # noqa: E402
from orb_slam_b200 import ORBextractor              # noqa: E402
from orb_slam_b200.extractor import OrbxError       # noqa: E402


def make_image(rng, kind, h, w, seed):
    if kind == "scene":
        return O.synth(h, w, seed)
    if kind == "noise":
        return rng.integers(0, 256, (h, w), dtype=np.uint8)
    if kind == "lowcontrast":
        return (O.synth(h, w, seed) // int(rng.integers(4, 24)) + int(rng.integers(0, 200))).astype(np.uint8)
    b = int(rng.integers(3, 40))                    # blocks: hard edges and corners on a coarse lattice
    yy, xx = np.mgrid[0:h, 0:w]
    base = (((yy // b) * 37 + (xx // b) * 101) % 256).astype(np.uint8)
    return base ^ (rng.integers(0, 4, (h, w), dtype=np.uint8))


def same(k, d, ok, od):
    if len(k) != len(ok):
        return f"count {len(k)} vs {len(ok)}"
    for f in ("octave", "x", "y", "response", "size", "class_id"):
        if not np.array_equal(k[f], ok[f]):
            return f"field {f}"
    if not np.array_equal(k["angle"].view(np.uint32), ok["angle"].view(np.uint32)):
        w = np.flatnonzero(k["angle"].view(np.uint32) != ok["angle"].view(np.uint32))
        i = w[0]
        return (f"angle bits at {len(w)} of {len(k)} keypoints; first: octave {k['octave'][i]} x {k['x'][i]} y {k['y'][i]} "
                f"angle {k['angle'][i]!r} vs {ok['angle'][i]!r}; octaves hit {sorted(set(k['octave'][w].tolist()))}")
    if not np.array_equal(d, od):
        return "descriptor bytes"
    return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=200)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--only", type=str, default="", help="comma-separated case numbers to run (the stream stays the same)")
    a = ap.parse_args()
    only = {int(x) for x in a.only.split(",") if x}
    rng = np.random.default_rng(a.seed)
    bad = degenerate = kp_total = tiny = big_cell = 0
    for c in range(a.cases):
        w, h = int(rng.integers(70, 1400)), int(rng.integers(70, 800))
        cfg = dict(nfeatures=int(rng.choice([50, 200, 500, 1000, 2000, 3000])),
                   scaleFactor=float(rng.choice([1.1, 1.2, 1.25, 1.3, 1.5, 2.0])), nlevels=int(rng.integers(1, 11)),
                   scoreType=int(rng.integers(0, 2)), fastTh=int(rng.integers(1, 45)))
        modes = dict(blur_mode=int(rng.integers(0, 2)), retain_mode=int(rng.integers(0, 2)))
        kind = str(rng.choice(["scene", "scene", "noise", "lowcontrast", "blocks"]))
        nfr = int(rng.integers(1, 4))
        imgs = np.stack([make_image(rng, kind, h, w, 1000 * c + i) for i in range(nfr)])
        tag = f"case {c}: {w}x{h} {kind} x{nfr} {cfg} {modes}"
        if only and c not in only:
            continue
        # a pyramid level that does not hold the 16-px border on both sides: the reference computes negative cell sizes
        # and asserts in cv::Mat; the CUDA path refuses the configuration at orbx_create, the oracle is not asked
        if min(w, h) / np.float32(cfg["scaleFactor"]) ** (cfg["nlevels"] - 1) < 40:
            tiny += 1
            continue
        ref = None
        try:
            oc = O.OracleExtractor(**cfg, **modes)
            ref = [oc(im) for im in imgs]
        except RuntimeError:
            degenerate += 1
        try:
            ex = ORBextractor(**cfg, **modes, max_width=w, max_height=h, max_batch=2)
            got = [ex(im) for im in imgs] if nfr == 1 else None
            if got is None:
                kps, desc, n = ex.extract_batch(imgs)
                got = [(kps[i, :n[i]], desc[i, :n[i]]) for i in range(nfr)]
        except OrbxError as e:
            if "cell too large" in str(e):          # stated limit of the CUDA path: refused loudly
                big_cell += 1
                continue
            if ref is not None:
                bad += 1
                print(f"FAIL {tag}: CUDA path refused ({e}) what the oracle accepts", flush=True)
            continue
        if ref is None:
            bad += 1
            print(f"FAIL {tag}: oracle degenerate, CUDA path ran", flush=True)
            continue
        for i, ((k, d), (ok, od)) in enumerate(zip(got, ref)):
            why = same(k, d, ok, od)
            kp_total += len(ok)
            if why:
                bad += 1
                print(f"FAIL {tag}: frame {i}: {why}", flush=True)
                break
    print(f"{a.cases} cases, {tiny} skipped (a level smaller than the border), {degenerate} degenerate (refused by both), "
          f"{big_cell} refused for a cell beyond the shared-memory FAST tile, {kp_total} keypoints compared, {bad} failures")
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
