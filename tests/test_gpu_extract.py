"""Parity of the CUDA extractor (through the C ABI, include/orbx.h) with the CPU oracle, stage by stage and end to
end.  Bit-exact everywhere: coordinates, octaves, responses, angles (raw bits), sizes, descriptors."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _ex(**kw):
    from orb_slam_b200 import ORBextractor
    return ORBextractor(**kw)


def _cmp_final(k, d, ok, od):
    assert len(k) == len(ok)
    for f in ("octave", "x", "y", "response", "size", "class_id"):
        assert np.array_equal(k[f], ok[f]), f
    assert np.array_equal(k["angle"].view(np.uint32), ok["angle"].view(np.uint32))  # raw bits (contract: 1e-3 deg)
    assert np.array_equal(d, od)


CONFIGS = [
    dict(w=640, h=480, nf=1000, nl=8, th=20, seed=0),     # BASELINE config 1
    dict(w=1241, h=376, nf=2000, nl=8, th=20, seed=3),    # BASELINE config 3 (wide-image grid)
    dict(w=640, h=480, nf=2000, nl=8, th=20, seed=1),     # the initialisation extractor (src/Tracking.cc:128)
    dict(w=320, h=240, nf=500, nl=6, th=12, seed=2),
    dict(w=333, h=251, nf=700, nl=5, th=5, seed=4),       # odd sizes, fastTh below the retry threshold
]


@pytest.mark.parametrize("cfg", CONFIGS)
def test_stagewise_parity(oracle, cfg):
    img = oracle.synth(cfg["h"], cfg["w"], cfg["seed"])
    O = oracle.OracleExtractor(cfg["nf"], 1.2, cfg["nl"], 1, cfg["th"])
    ok, od = O(img)
    ex = _ex(nfeatures=cfg["nf"], scaleFactor=1.2, nlevels=cfg["nl"], scoreType=1, fastTh=cfg["th"],
             max_width=cfg["w"], max_height=cfg["h"], max_batch=2)
    k, d = ex(img)
    for l in range(cfg["nl"]):
        info = ex.debug_level_info(l)
        g = O.level_grid(l)
        assert (info["w"], info["h"]) == O.level_size(l)
        assert (info["cols"], info["rows"], info["cellW"], info["cellH"], info["quota"]) == (
            g["cols"], g["rows"], g["cellW"], g["cellH"], g["quota"])
        assert np.array_equal(ex.debug_level_image(0, l, 0), O.level_image(l, 0)), f"pyramid level {l}"
        if g["nkept"]:
            assert np.array_equal(ex.debug_level_image(0, l, 1), O.level_image(l, 1)), f"blur level {l}"
        nt, nr = ex.debug_level_cells(0, l)
        ont, onr = O.level_cells(l)
        assert np.array_equal(nt, ont), f"nTotal level {l}"
        assert np.array_equal(nr, onr), f"nToRetain level {l}"
    _cmp_final(k, d, ok, od)


@pytest.mark.parametrize("w,h,nf,sf,nl,st,th,seed", [
    (459, 193, 2000, 1.2, 6, 0, 3, 21000),      # level 5 (184x78): 9 rows of ceil(46 / 9) = 6 px overshoot H = 46
    (459, 193, 2000, 1.2, 6, 1, 20, 5),
    (1331, 79, 500, 1.5, 2, 1, 20, 7),          # ratio 16.8: 33 one-pixel rows over H = 21
])
def test_cell_grid_overshooting_the_border(oracle, w, h, nf, sf, nl, st, th, seed):
    """ceil() in the cell size (:539-540) lets the rows before the last one run past maxBorderY: corners closer than
    15 px to the bottom edge, whose IC_Angle disc and BRIEF pattern read the reference's reflect-101 frame (:806-814).
    Found by tools/fuzz_extract_parity.py."""
    img = oracle.synth(h, w, seed)
    O = oracle.OracleExtractor(nf, sf, nl, st, th)
    ok, od = O(img)
    scale = O.tables()[1]
    near = 0
    for l in range(nl):
        lw, lh = O.level_size(l)
        m = ok["octave"] == l
        near += int(((np.rint(ok["x"][m] / scale[l]) > lw - 16) | (np.rint(ok["y"][m] / scale[l]) > lh - 16)).sum())
    assert near > 0, "the scene no longer exercises the overshoot"
    ex = _ex(nfeatures=nf, scaleFactor=sf, nlevels=nl, scoreType=st, fastTh=th, max_width=w, max_height=h, max_batch=2)
    k, d = ex(img)
    _cmp_final(k, d, ok, od)
    kps, desc, n = ex.extract_batch(np.stack([img, img, img]))
    for i in range(3):
        _cmp_final(kps[i, :n[i]], desc[i, :n[i]], ok, od)


def test_cell_candidates(oracle):
    """Per-cell FAST candidates (after NMS and the th=7 retry rule) equal cv::FAST's per-cell output as a set; their
    row-major ORDER is what std::nth_element later depends on and is covered by the end-to-end order parity."""
    img = oracle.synth(480, 640, 9)
    O = oracle.OracleExtractor()
    O(img)
    ex = _ex(max_batch=1)
    ex(img)
    ncheck = 0
    for l in range(8):
        g = O.level_grid(l)
        for c in range(g["cols"] * g["rows"]):
            ox, oy, orr = O.cell_keys(l, c)
            x, y, r = ex.debug_cell_candidates(0, l, c)
            j, i = c % g["cols"], c // g["cols"]
            ix, iy = 16 + j * g["cellW"] - 3, 16 + i * g["cellH"] - 3
            assert sorted(zip(oy.tolist(), ox.tolist(), orr.tolist())) == sorted(
                zip((y - iy).tolist(), (x - ix).tolist(), r.tolist()))
            ncheck += len(x)
    assert ncheck > 2000


@pytest.mark.parametrize("blur_mode,retain_mode,score", [(1, 0, 1), (0, 1, 1), (1, 1, 1), (0, 0, 0), (1, 1, 0)])
def test_modes(oracle, blur_mode, retain_mode, score):
    img = oracle.synth(480, 640, 21)
    ok, od = oracle.OracleExtractor(1000, 1.2, 8, score, 20, blur_mode, retain_mode)(img)
    k, d = _ex(nfeatures=1000, scoreType=score, blur_mode=blur_mode, retain_mode=retain_mode, max_batch=1)(img)
    _cmp_final(k, d, ok, od)


def test_texture_extremes(oracle):
    """Noise (thousands of corners per cell -> deep nth_element recursion), flat (zero keypoints, th=7 retry
    everywhere) and half/half images."""
    rng = np.random.default_rng(5)
    noise = rng.integers(0, 256, (480, 640), dtype=np.uint8)
    flat = np.full((480, 640), 77, np.uint8)
    half = flat.copy(); half[:, 320:] = noise[:, 320:]
    ex = _ex(max_batch=1)
    O = oracle.OracleExtractor()
    for img in (noise, flat, half):
        ok, od = O(img)
        k, d = ex(img)
        _cmp_final(k, d, ok, od)
    assert len(ex(flat)[0]) == 0


def test_batch_and_strides(oracle):
    """Batch path (chunking across max_batch, pipelined copies) == per-frame oracle; strided input rows."""
    F, H, W = 7, 240, 320
    frames = np.stack([oracle.synth(H, W, 100 + s) for s in range(F)])
    ex = _ex(nfeatures=500, nlevels=6, max_width=W, max_height=H, max_batch=3)
    O = oracle.OracleExtractor(500, 1.2, 6, 1, 20)
    kps, desc, n = ex.extract_batch(frames)
    for f in range(F):
        ok, od = O(frames[f])
        _cmp_final(kps[f, :n[f]], desc[f, :n[f]], ok, od)
    padded = np.zeros((H, W + 37), np.uint8)
    padded[:, :W] = frames[2]
    k, d = ex(padded[:, :W])
    ok, od = O(frames[2])
    _cmp_final(k, d, ok, od)


def test_empty_and_errors():
    from orb_slam_b200 import ORBextractor, OrbxError
    ex = ORBextractor(max_batch=1)
    k, d = ex(np.zeros((0, 0), np.uint8))
    assert len(k) == 0 and d.shape == (0, 32)
    with pytest.raises(OrbxError):
        ex(np.zeros((481, 640), np.uint8))          # larger than max_height
    with pytest.raises(OrbxError):
        ORBextractor(nfeatures=50, max_batch=1)     # zero-column grid: the reference divides by zero
    assert ex.GetLevels() == 8 and abs(ex.GetScaleFactor() - 1.2) < 1e-6


def test_synth_twin(oracle):
    import torch
    from orb_slam_b200 import synth_frames_device
    F, H, W = 3, 120, 200
    buf = torch.empty((F, H, W), dtype=torch.uint8, device="cuda")
    synth_frames_device(buf.data_ptr(), F, W, H, W, H * W, seed0=40)
    got = buf.cpu().numpy()
    for f in range(F):
        assert np.array_equal(got[f], oracle.synth(H, W, 40 + f))


def test_device_path_matches_host_path(oracle):
    import torch
    from orb_slam_b200 import KP_DTYPE
    F, H, W = 5, 240, 320
    frames = np.stack([oracle.synth(H, W, 300 + s) for s in range(F)])
    ex = _ex(nfeatures=500, nlevels=6, max_width=W, max_height=H, max_batch=2)
    kps, desc, n = ex.extract_batch(frames)
    d_img = torch.from_numpy(frames).cuda()
    d_k = torch.zeros((F, 500, 28), dtype=torch.uint8, device="cuda")
    d_d = torch.zeros((F, 500, 32), dtype=torch.uint8, device="cuda")
    d_n = torch.zeros(F, dtype=torch.int32, device="cuda")
    ex.extract_batch_device(d_img.data_ptr(), F, W, H, W, H * W, d_k.data_ptr(), d_d.data_ptr(), d_n.data_ptr(),
                            torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    ex.synchronize()
    n2 = d_n.cpu().numpy()
    assert np.array_equal(n2, n)
    k2 = d_k.cpu().numpy().view(KP_DTYPE).reshape(F, 500)
    for f in range(F):
        assert np.array_equal(k2[f, :n[f]], kps[f, :n[f]])
        assert np.array_equal(d_d[f, :n[f]].cpu().numpy(), desc[f, :n[f]])


def test_device_input_unaligned(oracle):
    """Caller-owned device frames with an odd base address and odd row stride: every kernel's unaligned path
    (resize slow path, FAST funnel-shift staging, blur byte staging, describe) must give the aligned result."""
    import torch
    from orb_slam_b200 import KP_DTYPE
    F, H, W = 3, 251, 333
    frames = np.stack([oracle.synth(H, W, 300 + s) for s in range(F)])
    ex = _ex(nfeatures=600, nlevels=6, max_width=W, max_height=H, max_batch=2)
    kps, desc, n = ex.extract_batch(frames)
    stride, fstride, base = W + 5, H * (W + 5) + 11, 3
    buf = torch.zeros(base + F * fstride + 64, dtype=torch.uint8, device="cuda")
    for f in range(F):
        view = buf[base + f * fstride: base + f * fstride + H * stride].view(H, stride)
        view[:, :W] = torch.from_numpy(frames[f]).cuda()
    d_k = torch.zeros((F, 600, 28), dtype=torch.uint8, device="cuda")
    d_d = torch.zeros((F, 600, 32), dtype=torch.uint8, device="cuda")
    d_n = torch.zeros(F, dtype=torch.int32, device="cuda")
    ex.extract_batch_device(buf.data_ptr() + base, F, W, H, stride, fstride, d_k.data_ptr(), d_d.data_ptr(),
                            d_n.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    ex.synchronize()
    assert np.array_equal(d_n.cpu().numpy(), n)
    k2 = d_k.cpu().numpy().view(KP_DTYPE).reshape(F, 600)
    for f in range(F):
        assert np.array_equal(k2[f, :n[f]], kps[f, :n[f]])
        assert np.array_equal(d_d[f, :n[f]].cpu().numpy(), desc[f, :n[f]])


@pytest.mark.parametrize("div,seed", [(3, 11), (6, 12)])
def test_low_contrast_retry_path(oracle, div, seed):
    """Low-contrast frames: most cells have <= 3 corners at fastTh = 20 and are searched again at 7 (:609-614) — the
    second pass of fast_cells_kernel — while a few keep their fastTh result; per-cell counts and the final output
    must match the oracle."""
    img = (oracle.synth(480, 640, seed).astype(np.int32) // div + 90).astype(np.uint8)
    O = oracle.OracleExtractor(1000, 1.2, 8, 1, 20)
    ok, od = O(img)
    ex = _ex(nfeatures=1000, nlevels=8, max_width=640, max_height=480, max_batch=1)
    k, d = ex(img)
    for l in range(8):
        nt, nr = ex.debug_level_cells(0, l)
        ont, onr = O.level_cells(l)
        assert np.array_equal(nt, ont), f"nTotal level {l}"
        assert np.array_equal(nr, onr), f"nToRetain level {l}"
    _cmp_final(k, d, ok, od)


def test_cuda_equals_reference_generated_golden():
    """The CUDA extractor against vectors produced by the reference's OWN src/ORBextractor.cc (compiled in place, OpenCV
    routines served by cv2 4.13; tests/golden/make_ref_extract_golden.py): every field and descriptor byte, for both
    retainBest variants, with the OpenCV >= 3.4 Gaussian taps (blur_mode = 1)."""
    import os
    from orb_slam_b200 import ORBextractor
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_extract.npz"))
    for c in range(int(g["ncases"])):
        nf, nl, st, th, sf = (int(v) for v in g[f"params{c}"])
        img = g[f"img{c}"]
        for mode in (1, 0):
            ex = ORBextractor(nf, sf / 1000.0, nl, st, th, max_width=img.shape[1], max_height=img.shape[0], max_batch=2,
                              blur_mode=1, retain_mode=mode)
            k, d = ex(img)
            rk, rd = g[f"kps{c}_{mode}"], g[f"desc{c}_{mode}"]
            assert len(k) == len(rk), (c, mode)
            for f in ("x", "y", "size", "angle", "response", "octave", "class_id"):
                a, b = k[f], rk[f]
                if a.dtype.kind == "f":
                    a, b = a.view(np.uint32), b.view(np.uint32)
                assert np.array_equal(a, b), (c, mode, f)
            assert np.array_equal(d, rd), (c, mode)
            ex.close()


@pytest.mark.parametrize("env", [
    {"ORBX_SELECT": "global", "ORBX_PDL": "0", "ORBX_LANES": "1"},
    {"ORBX_SELECT": "smem", "ORBX_PDL": "2", "ORBX_LANES": "3"},
    {"ORBX_SELECT": "global", "ORBX_PDL": "2", "ORBX_LANES": "4"},
    {"ORBX_SELECT": "smem", "ORBX_PDL": "0", "ORBX_LANES": "2"},
])
def test_runtime_switches_are_bit_identical(oracle, env):
    """The run-time switches choose between code paths that must give identical results: both selection
    kernels on single frames AND on a multi-chunk batch, programmatic dependent launch never / always, 1-4 chunk lanes.
    They are read when the handle is created."""
    import os
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        ex = _ex(nfeatures=700, nlevels=6, max_width=400, max_height=300, max_batch=8)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    O = oracle.OracleExtractor(700, 1.2, 6, 1, 20)
    imgs = np.stack([oracle.synth(300, 400, 50 + s) for s in range(5)])
    ref = [O(im) for im in imgs]
    k, d = ex(imgs[0])
    _cmp_final(k, d, *ref[0])
    batch = np.concatenate([imgs] * 5)                       # 25 frames = 4 chunks of <= 8 over the lanes
    kps, desc, n = ex.extract_batch(batch)
    for i in range(len(batch)):
        _cmp_final(kps[i, :n[i]], desc[i, :n[i]], *ref[i % 5])


def test_fast_kernels_agree(oracle):
    """The TMA-staged FAST / pyramid kernels (default where the buffers allow it) and the legacy kernels (ORBX_FAST=legacy,
    ORBX_RESIZE=legacy; also the fallback for misaligned caller buffers) on scenes that take every branch: list overflow (noise), retry at 7 (flat /
    low contrast), odd sizes whose cells straddle the tensor-copy alignment, a row pitch that is not a multiple of 16."""
    import os
    rng = np.random.default_rng(11)
    cases = [(640, 480, 1000, 8, 20, "scene"), (333, 251, 700, 5, 9, "noise"), (459, 193, 2000, 6, 20, "low"),
             (1241, 376, 2000, 8, 20, "half"), (401, 301, 300, 4, 5, "scene")]
    for w, h, nf, nl, th, kind in cases:
        base = oracle.synth(h, w, 77)
        img = {"scene": base, "noise": rng.integers(0, 256, (h, w), dtype=np.uint8),
               "low": (base // 9 + 60).astype(np.uint8),
               "half": np.where(np.arange(w)[None, :] < w // 2, 90, rng.integers(0, 256, (h, w))).astype(np.uint8)}[kind]
        ok, od = oracle.OracleExtractor(nf, 1.2, nl, 1, th)(img)
        for mode in ("tma", "legacy"):
            os.environ["ORBX_FAST"] = os.environ["ORBX_RESIZE"] = mode     # FAST and pyramid kernels alike
            try:
                ex = _ex(nfeatures=nf, nlevels=nl, fastTh=th, max_width=w, max_height=h, max_batch=3)
            finally:
                os.environ.pop("ORBX_FAST", None); os.environ.pop("ORBX_RESIZE", None)
            k, d = ex(img)
            _cmp_final(k, d, ok, od)
            kps, desc, n = ex.extract_batch(np.stack([img] * 5))       # 2 chunks, both lanes
            for i in range(5):
                _cmp_final(kps[i, :n[i]], desc[i, :n[i]], ok, od)


def test_device_api_with_unaligned_rows_takes_the_legacy_kernels(oracle):
    """orbx_extract_batch_device on a caller buffer whose row pitch (333) and frame stride are not multiples of 16: the
    tensor maps of the TMA kernels cannot describe it, so the chunk runs on the legacy kernels - same results."""
    import torch
    from orb_slam_b200 import KP_DTYPE
    w, h, nf = 333, 251, 700
    imgs = np.stack([oracle.synth(h, w, 90 + s) for s in range(3)])
    ex = _ex(nfeatures=nf, nlevels=5, fastTh=9, max_width=w, max_height=h, max_batch=4)
    d = torch.from_numpy(imgs).cuda()
    assert d.stride(1) == 333
    d_k = torch.zeros((3, nf, 28), dtype=torch.uint8, device="cuda")
    d_d = torch.zeros((3, nf, 32), dtype=torch.uint8, device="cuda")
    d_n = torch.zeros(3, dtype=torch.int32, device="cuda")
    ex.extract_batch_device(d.data_ptr(), 3, w, h, w, w * h, d_k.data_ptr(), d_d.data_ptr(), d_n.data_ptr(),
                            torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    ex.synchronize()
    n = d_n.cpu().numpy(); k = d_k.cpu().numpy().view(KP_DTYPE).reshape(3, nf); dd = d_d.cpu().numpy()
    O = oracle.OracleExtractor(nf, 1.2, 5, 1, 9)
    for f in range(3):
        ok, od = O(imgs[f])
        _cmp_final(k[f, :n[f]], dd[f, :n[f]], ok, od)


@pytest.mark.parametrize("w,h,nf,nl,st,th", [(975, 686, 200, 8, 1, 20), (1200, 700, 150, 4, 0, 12), (900, 640, 100, 3, 1, 3)])
def test_cells_beyond_shared_memory(oracle, w, h, nf, nl, st, th):
    """Few features on a large image give cells of > 60 k pixels (2 x 2 cells of 472 x 327 at 975 x 686 / 200 features):
    too large for the shared-memory FAST kernels, served by fast_cells_big_kernel in global memory.  Scene, noise (thousands
    of corners per cell) and low-contrast (retry at 7) images, FAST and Harris scores, mixed with normal cells on the
    smaller levels."""
    rng = np.random.default_rng(3)
    base = oracle.synth(h, w, 31)
    imgs = [base, rng.integers(0, 256, (h, w), dtype=np.uint8), (base // 12 + 90).astype(np.uint8)]
    O = oracle.OracleExtractor(nf, 1.2, nl, st, th)
    ex = _ex(nfeatures=nf, nlevels=nl, scoreType=st, fastTh=th, max_width=w, max_height=h, max_batch=2)
    big = 0
    for l in range(nl):
        i = ex.debug_level_info(l)
        big += int((i["cellW"] + 6) * (i["cellH"] + 6) > 60_000)       # ~4.5 B of shared memory per pixel against 200 KB
    assert big >= 1, "the configuration no longer has an oversized cell"
    ref = [O(im) for im in imgs]
    for im, (ok, od) in zip(imgs, ref):
        k, d = ex(im)
        _cmp_final(k, d, ok, od)
    kps, desc, n = ex.extract_batch(np.stack(imgs))
    for i in range(3):
        _cmp_final(kps[i, :n[i]], desc[i, :n[i]], *ref[i])
