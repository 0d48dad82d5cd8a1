"""The drop-in boundary as CODE: the reference's UNMODIFIED src/Frame.cc / KeyFrame.cc / MapPoint.cc compiled
against include/shim/ORBextractor.h + include/shim/ORBmatcher.h and linked with liborbx.so
(oracle/_ref/libref_dropin.so, `make -C oracle ref_dropin`).  Its Frame constructor runs `(*mpORBextractor)(im, cv::Mat(),
mvKeys, mDescriptors)` (src/Frame.cc:59) on the CUDA extractor, and the reference-signature matcher methods run on the
CUDA matcher.  Everything is compared with the SAME scenario through libref_orbslam.so, the all-CPU reference (its
src/ORBextractor.cc and src/ORBmatcher.cc): mvKeys, mvKeysUn, mDescriptors, the grid, and the results of
SearchForInitialization / WindowSearch / SearchByProjection(F1, F2) / SearchByBoW x2, all bit-identical.
Also: a plain-C translation unit that includes orbx.h (proves the header is C and the ABI is callable from C)."""
import os
import subprocess

import numpy as np
import pytest

import refscene as RS

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _rf():
    from oracle import ref_frontend as RF
    if not RF.available() or not os.path.exists(RF.DROPIN_LIB):
        pytest.skip("compiled reference front-end (oracle/_ref) / cv2 not available here")
    return RF


@pytest.fixture()
def cv4_modes(monkeypatch):
    # cv2 4.13 serves the CPU reference's primitives: ask the shim for that OpenCV's arithmetic (include/orbx.h)
    monkeypatch.setenv("ORBX_BLUR_MODE", "1")
    monkeypatch.setenv("ORBX_RETAIN_MODE", "1")


def _scenario(RF, oracle, w, h, nfeat, nlevels, cam, seed, vpath=None):
    """Everything observable of one two-frame scenario through the library RF currently points at."""
    world = RF.World(vpath, nfeat, 1.2, nlevels, 1, 20)
    ims = RS.images(oracle, seed, h, w)
    fs = [world.frame(ims[0], RS.K_of(cam), cam[4:9], reset_static=True), world.frame(ims[1], RS.K_of(cam), cam[4:9])]
    out = {}
    for k, f in enumerate(fs):
        out[f"f{k}"] = (f.kps.copy(), f.xu.copy(), f.yu.copy(), f.desc.copy(), f.bounds, f.inv, *f.grid())
    prev0 = np.stack([fs[0].xu, fs[0].yu], 1)
    out["sfi"] = [RF.search_for_initialization(fs[0], fs[1], prev0, win, ratio, ori)
                  for win, ratio, ori in ((100, 0.9, True), (30, 0.8, False))]
    v = [RS.valid_flags(f.n, seed + k) for k, f in enumerate(fs)]
    for k, f in enumerate(fs):
        f.set_valid(v[k])
    out["win"] = [RF.window_search(fs[0], fs[1], win, l0, l1, ratio, ori)
                  for win, l0, l1, ratio, ori in ((200, 0, 2**31 - 1, 0.8, True), (100, 1, 3, 0.8, False))]
    if vpath:
        out["bow_vec"] = [f.compute_bow() for f in fs]
        out["bow"] = [RF.search_by_bow(fs[0], fs[1], variant, ratio, ori)
                      for variant in ("kf_f", "kf_kf") for ratio, ori in ((0.75, True), (0.6, False))]
    rng = np.random.default_rng(seed)
    sf = np.cumprod(np.concatenate([[1.0], np.full(nlevels - 1, np.float32(1.2))]).astype(np.float32)).astype(np.float32)
    pos = RS._point_cloud(rng, fs[0], -11, -6, cam, sf, nlevels)[0]
    out["proj_ff"] = []
    for window, ratio in ((15, 0.8), (50, 0.9)):
        fs[0].set_valid(v[0]); fs[1].set_valid(v[1])
        out["proj_ff"].append(RF.search_by_projection_ff(fs[0], fs[1], pos, window, ratio, True))
    return out


def _same(a, b, path=""):
    if isinstance(a, (tuple, list)):
        assert len(a) == len(b), path
        for i, (x, y) in enumerate(zip(a, b)):
            _same(x, y, f"{path}[{i}]")
    elif isinstance(a, np.ndarray):
        assert a.dtype == b.dtype and a.shape == b.shape, path
        assert a.tobytes() == b.tobytes(), path          # raw bytes: float fields compared as bits
    else:
        assert a == b, path


@pytest.mark.parametrize("w,h,nfeat,nlevels,cam,seed", [
    (640, 480, 1000, 8, (537.9, 539.97, 315.2, 229.3, -0.4157, 0.2624, 0.0, 0.0, -0.1178), 61),   # config 1 / 2 shape
    (1241, 376, 2000, 8, (718.856, 718.856, 607.19, 185.2, 0.0, 0.0, 0.0, 0.0, 0.0), 62),          # config 3 (KITTI)
    (320, 240, 500, 6, RS.CAM_SHIPPED, 63),
])
def test_reference_frame_runs_on_the_cuda_extractor_and_matcher(oracle, tmp_path, cv4_modes, w, h, nfeat, nlevels, cam, seed):
    RF = _rf()
    from oracle import bow_oracle as B
    V = B.synth_vocabulary(k=3, L=5, seed=seed, stop_frac=0.02)
    vpath = str(tmp_path / "voc.txt")
    B.save_text(vpath, **V)
    cpu = _scenario(RF, oracle, w, h, nfeat, nlevels, cam, seed, vpath)
    from orb_slam_b200._lib import lib
    L = lib()                                              # the process-wide library the drop-in links against
    with RF.library(RF.DROPIN_LIB):
        gpu = _scenario(RF, oracle, w, h, nfeat, nlevels, cam, seed, vpath)
    assert len(cpu["f0"][0]) > 0.9 * nfeat
    assert sum(r[0] for r in cpu["sfi"]) > 50 and sum(r[0] for r in cpu["win"]) > 50
    assert sum(r[0] for r in cpu["bow"]) > 50 and sum(r[0] for r in cpu["proj_ff"]) > 50
    for key in cpu:
        _same(cpu[key], gpu[key], key)


C_PROGRAM = r"""
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "orbx.h"
int main(int argc, char** argv) {
  const int W = 320, H = 240;
  unsigned char* img = (unsigned char*)malloc((size_t)W * H);
  FILE* f = fopen(argv[1], "rb");
  if (!f || fread(img, 1, (size_t)W * H, f) != (size_t)W * H) return 2;
  fclose(f);
  orbx_params p;
  orbx_default_params(&p);
  p.nfeatures = 500; p.nlevels = 6; p.max_width = W; p.max_height = H; p.max_batch = 1;
  orbx_handle* h = NULL;
  if (orbx_create(&p, &h) != ORBX_OK) { printf("create: %s\n", orbx_last_error(NULL)); return 3; }
  orbx_keypoint* kp = (orbx_keypoint*)malloc(sizeof(orbx_keypoint) * 500);
  unsigned char* desc = (unsigned char*)malloc(500 * 32);
  int n = 0;
  if (orbx_extract(h, img, W, H, (size_t)W, kp, desc, &n) != ORBX_OK) { printf("extract: %s\n", orbx_last_error(h)); return 4; }
  unsigned long long sum = 1469598103934665603ull;        /* FNV-1a over keypoint records and descriptors */
  const unsigned char* b = (const unsigned char*)kp;
  for (size_t i = 0; i < sizeof(orbx_keypoint) * (size_t)n; i++) { sum ^= b[i]; sum *= 1099511628211ull; }
  for (size_t i = 0; i < (size_t)n * 32; i++) { sum ^= desc[i]; sum *= 1099511628211ull; }
  printf("%d %llu levels=%d scale=%.3f\n", n, sum, orbx_get_levels(h), orbx_get_scale_factor(h));
  orbx_destroy(h);
  return 0;
}
"""


def _fnv(*arrays):
    s = 1469598103934665603
    for a in arrays:
        for byte in a.tobytes():
            s = ((s ^ byte) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return s


def test_plain_c_program_calls_the_abi(oracle, tmp_path):
    """orbx.h from a C99 translation unit (gcc -std=c99 -pedantic-errors), linked with liborbx.so, one frame extracted;
    its checksum equals the oracle's keypoints + descriptors."""
    src = tmp_path / "abi.c"; exe = tmp_path / "abi"
    src.write_text(C_PROGRAM)
    libdir = os.path.join(ROOT, "orb_slam_b200")
    subprocess.run(["gcc", "-std=c99", "-pedantic-errors", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                    str(src), "-o", str(exe), "-L", libdir, "-lorbx", f"-Wl,-rpath,{libdir}"], check=True)
    img = oracle.synth(240, 320, 5)
    (tmp_path / "img.bin").write_bytes(img.tobytes())
    r = subprocess.run([str(exe), str(tmp_path / "img.bin")], capture_output=True, text=True, check=True)
    n, checksum = r.stdout.split()[:2]
    ok, od = oracle.OracleExtractor(500, 1.2, 6, 1, 20)(img)
    assert int(n) == len(ok) and int(checksum) == _fnv(ok, od), r.stdout
    assert "levels=6 scale=1.200" in r.stdout
