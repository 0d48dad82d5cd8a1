"""In-tree build of liborbx.so (hand-written sm_100a CUDA + the C-ABI of include/orbx.h, include/orbm.h).

`python -m orb_slam_b200.build` or `__graft_entry__.build()`.  nvcc cross-compiles without a GPU; the built .so is
git-ignored but travels to the GPU box with the repo snapshot.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "liborbx.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-Xcompiler", "-fPIC",
         "-shared", "-cudart", "static"]


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _deps():
    d = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    inc = os.path.join(HERE, "..", "include")
    d += [os.path.join(inc, f) for f in os.listdir(inc)]
    return d


def is_stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(p) > t for p in _deps())


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not is_stale():
        return LIB
    # experiments: ORBX_NVCC_EXTRA="-DORBX_FAST_THREADS=128" ORBX_LIB_OUT=/path/variant.so python -m orb_slam_b200.build --force
    out = os.environ.get("ORBX_LIB_OUT", LIB)
    cmd = [NVCC, *FLAGS, *os.environ.get("ORBX_NVCC_EXTRA", "").split(), "-o", out, *sources()]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
        print(" ".join(cmd))
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
