"""Matching legs of bench.py (BASELINE.json metric, second half: "Hamming matches/s").

config 5  all-pairs 2-NN: 5,000 keyframes x 1,000 descriptors (5 M x 256 bit), database sharded by keyframe over the
          ranks (STRONG scaling: the database is fixed), queries replicated, one ncclAllGather of the packed per-shard
          top-2 keys + merge kernel per query batch (orbm_knn2_sharded).
config 4  SearchForInitialization (window 100, ratio 0.9, orientation check) on 1000-keypoint frame pairs cut from
          shifted synthetic images and extracted by the CUDA extractor; batched over all pairs.
Returns a dict that bench.py embeds under "matching" in its single JSON line.
"""
from __future__ import annotations

import os
import time

import numpy as np

DB_KEYFRAMES, KP_PER_KF = 5000, 1000
# 8 x XOR32 + 8 x POPC32 per descriptor pair; POPC issues 16 lanes/clk/SM on sm_100
# (tools/popc_microbench.cu, measured on this pool's B200: 0.50 warp-POPC/clk/SM)
POPC_PER_CLK_PER_SM = 16.0


def bench_knn2(args, rank, world, dev, matcher, clocks_mhz):
    import torch
    import torch.distributed as dist
    from orb_slam_b200 import ShardedKnn2, shard_range

    kf0, kf1 = shard_range(DB_KEYFRAMES, rank, world)
    ndb = (kf1 - kf0) * KP_PER_KF
    g = torch.Generator(device=dev); g.manual_seed(12345 + rank)
    db = torch.randint(0, 256, (ndb, 32), dtype=torch.uint8, device=dev, generator=g)
    nq = args.query_keyframes * KP_PER_KF
    # queries (SURVEY config 5): database descriptors + bit-flip noise (each bit w.p. 0.1), so true matches exist and
    # the accept branch of the merge runs.  Drawn from the first keyframes of rank 0's shard (database rows 0 .. nq),
    # replicated on every rank.
    gq = torch.Generator(device=dev); gq.manual_seed(999)
    q = torch.empty((nq, 32), dtype=torch.uint8, device=dev)
    if rank == 0:
        flips = torch.zeros((nq, 32), dtype=torch.uint8, device=dev)
        for b in range(8):
            flips |= (torch.rand((nq, 32), device=dev, generator=gq) < 0.1).to(torch.uint8) << b
        q.copy_(db[:nq] ^ flips)
    if world > 1:
        dist.broadcast(q, 0)
    gathered = torch.empty((world, nq, 2), dtype=torch.int32, device=dev)
    top2 = torch.empty((nq, 2), dtype=torch.int32, device=dev)
    match = torch.empty(nq, dtype=torch.int32, device=dev)
    stream = torch.cuda.Stream(device=dev)
    sh = ShardedKnn2.from_torch_distributed(matcher, dev) if world > 1 else None

    def step():
        if sh is not None:
            sh.search(q.data_ptr(), nq, db.data_ptr(), ndb, kf0 * KP_PER_KF, gathered.data_ptr(), top2.data_ptr(),
                      match.data_ptr(), 50, 0.75, stream.cuda_stream)
        else:
            matcher.knn2_device(q.data_ptr(), nq, db.data_ptr(), ndb, 0, gathered.data_ptr(), stream.cuda_stream)
            matcher.merge_top2_device(gathered.data_ptr(), 1, nq, 50, 0.75, top2.data_ptr(), match.data_ptr(),
                                      stream.cuda_stream)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    l0 = matcher.kernel_launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1) / args.steps
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    launches = matcher.kernel_launches() - l0
    # ---- result check (outside the timed region): the sharded search (local 2-NN -> ncclAllGather -> merge) against ONE
    # single-rank search over the whole database, which rank 0 regenerates from the ranks' seeds, on a query sample
    check = None
    if rank == 0:
        ns = min(nq, 512)
        ref = torch.empty((ns, 2), dtype=torch.int32, device=dev)
        if world > 1:
            parts = []
            for r in range(world):
                a, b = shard_range(DB_KEYFRAMES, r, world)
                gr = torch.Generator(device=dev); gr.manual_seed(12345 + r)
                parts.append(torch.randint(0, 256, ((b - a) * KP_PER_KF, 32), dtype=torch.uint8, device=dev, generator=gr))
            full = torch.cat(parts); del parts
        else:
            full = db
        matcher.knn2_device(q.data_ptr(), ns, full.data_ptr(), full.shape[0], 0, ref.data_ptr(), stream.cuda_stream)
        torch.cuda.synchronize()
        same = bool(torch.equal(ref, top2[:ns]))
        best = ref[:, 0].to(torch.int64) & 0xFFFFFFFF
        true_hit = float(((best & 0x7FFFFF) == torch.arange(ns, device=dev)).float().mean().item())
        check = {"queries_checked": ns, "sharded_equals_single": same, "true_match_is_best_frac": true_hit,
                 "accepted_frac": float((match >= 0).float().mean().item())}
        if world > 1:
            del full
    pairs = float(nq) * DB_KEYFRAMES * KP_PER_KF
    pps = pairs / (ms * 1e-3)
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    mhz = clocks_mhz or 1965.0
    popc_peak = sms * POPC_PER_CLK_PER_SM * mhz * 1e6 * world
    out = {"metric": "Hamming distance evaluations/s (all-pairs 2-NN, 5k keyframes x 1000 descriptors)",
           "value": pps, "unit": "pairs/s", "ms_per_step": ms, "queries_per_step": nq, "db_descriptors": DB_KEYFRAMES * KP_PER_KF,
           "scaling": "strong", "shard_descriptors_per_gpu": ndb, "gpu_launches": launches,
           "collective": "ncclAllGather of [nq,2] u32 keys per step" if world > 1 else "none (1 GPU)",
           "roofline": knn2_roofline(pps, popc_peak, mhz, world),
           "db_bytes_read_per_step": float(ndb) * 32 * ((nq + 1023) // 1024), "check": check}
    if sh is not None:
        sh.close()
    return out


def knn2_roofline(pps, popc_peak, mhz, world):
    """M8 runs on the tensor cores (knn2_tc_kernel): 256 multiply-adds of +-1 FP8 values per descriptor pair = 512 flop.
    Peak = twice the dense bf16 throughput the driver measured on this pool (MEASURED_PEAKS.json, burst figure; FP8 issues at
    twice the bf16 rate), else twice the 2250 TFLOP/s nominal.  The former bound - 8 POPC32 per pair at 16 lanes/clk/SM -
    is kept beside it: a fraction above 1 there is the point of the tensor-core form."""
    import json, os
    bf16, src = 2250.0, "2 x nominal dense bf16 (2250 TFLOP/s): MEASURED_PEAKS.json absent"
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "MEASURED_PEAKS.json")) as f:
            bf16 = float(json.load(f)["bf16_tflops"]); src = "2 x MEASURED_PEAKS.json bf16_tflops (burst): fp8 issues at twice the bf16 rate"
    except Exception:
        pass
    peak = 2.0 * bf16 * world
    ach = pps * 512 / 1e12
    return {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak, "traffic": None,
            "dtype": "fp8 e4m3 operands (+-1), fp32 accumulate: exact integers", "peak_source": src,
            "popc_bound": {"achieved": pps * 8 / 1e12, "peak": popc_peak / 1e12, "frac": pps * 8 / popc_peak,
                           "unit": "T popc32/s (8 per pair): the bound of the POPC kernels this replaces",
                           "clock_mhz_assumed": mhz}}


def make_pairs(npairs, dev, local):
    """Frame pairs for config 4: two 640x480 windows of one 704x544 synthetic image, offset by (dx, dy)."""
    import torch
    from orb_slam_b200 import MatchFrame, ORBextractor, synth_frames_device
    W, H, BW, BH = 640, 480, 704, 544
    big = torch.empty((npairs, BH, BW), dtype=torch.uint8, device=dev)
    synth_frames_device(big.data_ptr(), npairs, BW, BH, BW, BH * BW, seed0=50_000)
    torch.cuda.synchronize()
    rng = np.random.default_rng(4)
    sh = rng.integers(-30, 31, (npairs, 2))
    a = big[:, 32:32 + H, 32:32 + W].contiguous()
    b = torch.stack([big[p, 32 + sh[p, 1]:32 + sh[p, 1] + H, 32 + sh[p, 0]:32 + sh[p, 0] + W] for p in range(npairs)]).contiguous()
    ex = ORBextractor(1000, 1.2, 8, 1, 20, max_width=W, max_height=H, max_batch=min(npairs, 128), device=local)
    ka, da, na = ex.extract_batch(a.cpu().numpy())
    kb, db, nb = ex.extract_batch(b.cpu().numpy())
    f1 = [MatchFrame.from_keypoints(ka[p, :na[p]], da[p, :na[p]], W, H) for p in range(npairs)]
    f2 = [MatchFrame.from_keypoints(kb[p, :nb[p]], db[p, :nb[p]], W, H) for p in range(npairs)]
    ex.close()
    return f1, f2


def bench_init_search(args, rank, world, dev, local, matcher):
    import torch
    import torch.distributed as dist
    from orb_slam_b200 import ORBmatcher
    npairs = args.match_pairs
    f1, f2 = make_pairs(npairs, dev, local)
    om = ORBmatcher(0.9, True, matcher=matcher)
    s1, s2 = om.upload(f1), om.upload(f2)
    idx = np.arange(npairs, dtype=np.int32)
    prev0 = [np.stack([f.x, f.y], 1).astype(np.float32) for f in f1]  # vbPrevMatched = F1 keypoints (Tracking.cc:334-336)
    stride1 = max(f.n for f in f1)
    prev_packed0 = np.zeros((npairs, stride1, 2), np.float32)
    for p in range(npairs):
        prev_packed0[p, :f1[p].n] = prev0[p]
    # host buffers of the e2e leg are pinned (torch), like bench.py's extraction leg
    prev_packed = torch.empty((npairs, stride1, 2), dtype=torch.float32, pin_memory=True).numpy()
    out_rows = torch.empty((npairs, stride1), dtype=torch.int32, pin_memory=True).numpy()
    out_nm = torch.empty(npairs, dtype=torch.int32, pin_memory=True).numpy()

    def step():
        np.copyto(prev_packed, prev_packed0)          # the search updates vbPrevMatched in place
        return om.SearchForInitialization(s1, s2, idx, idx, prev_packed, 100, out=out_rows, nm=out_nm)

    for _ in range(3):
        nm, rows = step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t = time.perf_counter(); kms = 0.0
    for _ in range(args.steps):
        nm, rows = step()
        kms += matcher.last_kernel_ms()
    e2e_ms = (time.perf_counter() - t) * 1e3 / args.steps
    kms /= args.steps
    if world > 1:
        tt = torch.tensor([e2e_ms, kms], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_ms, kms = float(tt[0].item()), float(tt[1].item())
    out = {"metric": "SearchForInitialization frame pairs/s (1000-kp pairs, window 100, ratio 0.9)",
           "value": npairs * world / (kms * 1e-3), "unit": "pairs/s", "kernel_ms_per_step": kms,
           "e2e": {"value": npairs * world / (e2e_ms * 1e-3), "unit": "pairs/s", "ms_per_step": e2e_ms,
                   "note": "pinned host prev_xy in, matches out through the C ABI; frame sets resident"},
           "pairs_per_gpu": npairs, "mean_matches_per_pair": float(nm.mean()), "scaling": "weak"}
    if rank == 0 and not args.no_cpu:
        from oracle import match_oracle as MO
        import ctypes as C
        L = MO.lib(True)
        n = min(npairs, 64)
        h1 = [MO.frame_struct(f) for f in f1[:n]]; h2 = [MO.frame_struct(f) for f in f2[:n]]
        A1 = (MO._Frame * n)(*[h.s for h in h1]); A2 = (MO._Frame * n)(*[h.s for h in h2])
        stride = max(f.n for f in f1[:n])
        threads = os.cpu_count() or 1

        def cpu(nthreads):
            xy = np.zeros((n, stride, 2), np.float32)
            for p in range(n):
                xy[p, :f1[p].n] = prev0[p]
            m = np.empty((n, stride), np.int32); cnt = np.empty(n, np.int32)
            t0 = time.perf_counter()
            L.mo_search_for_initialization_many(A1, A2, n, xy.ctypes.data, stride, 100, 0.9, 1, m.ctypes.data,
                                                cnt.ctypes.data, nthreads)
            return n / (time.perf_counter() - t0), m, cnt

        cpu(threads)
        v, m, cnt = cpu(threads)
        v1, _, _ = cpu(1)
        ok = all(np.array_equal(rows[p, :f1[p].n], m[p, :f1[p].n]) and nm[p] == cnt[p] for p in range(n))
        out["cpu_baseline"] = {"value": v, "unit": "pairs/s", "cores": threads, "kind": "port", "single_thread_value": v1,
                               "sample": f"{n} of the same pairs, oracle/liborb_oracle_fast.so", "gpu_equals_cpu": bool(ok)}
    return out


def bench_knn2_cpu(args):
    """Bounded CPU sample of config 5: 64 queries against a 200k-descriptor slice, all host threads."""
    from oracle import match_oracle as MO
    rng = np.random.default_rng(12345)
    db = rng.integers(0, 256, (200_000, 32), dtype=np.uint8)
    threads = os.cpu_count() or 1
    q = rng.integers(0, 256, (max(64, 4 * threads), 32), dtype=np.uint8)
    MO.knn2(q[:threads], db[:1000], fast=True, threads=threads)
    t = time.perf_counter()
    MO.knn2(q, db, fast=True, threads=threads)
    dt = time.perf_counter() - t
    t1 = time.perf_counter()
    MO.knn2(q[:4], db, fast=True, threads=1)
    dt1 = time.perf_counter() - t1
    return {"value": len(q) * len(db) / dt, "unit": "pairs/s", "cores": threads, "kind": "port",
            "single_thread_value": 4 * len(db) / dt1,
            "sample": f"{len(q)} queries x 200,000 database descriptors, oracle SWAR popcount (-O3 -march=native)"}


def bench_frame_glue(args, dev, local, matcher):
    """N1: extractor DEVICE output -> searchable frame set (undistort + 64x48 grid), the Frame constructor's
    glue (reference src/Frame.cc:59-127) with the shipped camera model scaled to 640x480.  CPU column: oracle
    undistortion + grid binning of the same keypoints, one host thread (what the reference's Tracking thread does)."""
    import torch
    from orb_slam_b200 import Camera, FrameSet, KP_DTYPE, ORBextractor, synth_frames_device
    W, H, NF, F = 640, 480, 1000, 256
    cam = (268.9633 * 2, 269.9858 * 2, 157.6087 * 2, 114.6369 * 2, -0.4157, 0.2624, 0.0, 0.0, -0.1178, W, H)
    d_frames = torch.empty((F, H, W), dtype=torch.uint8, device=dev)
    synth_frames_device(d_frames.data_ptr(), F, W, H, W, H * W, seed0=90_000)
    ex = ORBextractor(NF, 1.2, 8, 1, 20, max_width=W, max_height=H, max_batch=F, device=local)
    d_k = torch.zeros((F, NF, 28), dtype=torch.uint8, device=dev)
    d_d = torch.zeros((F, NF, 32), dtype=torch.uint8, device=dev)
    d_n = torch.zeros(F, dtype=torch.int32, device=dev)
    stream = torch.cuda.Stream(device=dev)
    ex.extract_batch_device(d_frames.data_ptr(), F, W, H, W, H * W, d_k.data_ptr(), d_d.data_ptr(), d_n.data_ptr(),
                            stream.cuda_stream)
    torch.cuda.synchronize()
    camera = Camera(*cam)

    def step():
        fs = FrameSet.from_extractor(matcher, d_k.data_ptr(), d_d.data_ptr(), d_n.data_ptr(), F, NF, camera,
                                     stream.cuda_stream)
        fs.close()
    for _ in range(3):
        step()
    t = time.perf_counter()
    for _ in range(max(args.steps, 5)):
        step()
    ms = (time.perf_counter() - t) * 1e3 / max(args.steps, 5)
    out = {"metric": "Frame glue frames/s (1000 kp: undistort + 64x48 grid, extractor device output -> frame set)",
           "value": F / (ms * 1e-3), "unit": "frames/s", "ms_per_step": ms, "frames_per_step": F,
           "note": "wall clock of orbm_frameset_from_extractor incl. its allocations, the 4 B/frame count read-back and "
                   "the final sync; features never leave the device"}
    if not args.no_cpu:
        from oracle import frame_oracle as FO
        from oracle import match_oracle as MO
        from orb_slam_b200 import MatchFrame
        kps = d_k.cpu().numpy().view(KP_DTYPE).reshape(F, NF)
        desc = d_d.cpu().numpy(); n = d_n.cpu().numpy()
        ocam = FO.Camera(*cam[:9])
        (mnx, mxx, mny, mxy), _ = FO.image_bounds(ocam, W, H)
        nc = 32
        t = time.perf_counter()
        for f in range(nc):
            k = kps[f, :n[f]]
            un = FO.undistort_points(ocam, np.stack([k["x"], k["y"]], 1))
            MO.grid_csr(MatchFrame(un[:, 0], un[:, 1], k["angle"], k["octave"], desc[f, :n[f]], min_x=float(mnx),
                                   min_y=float(mny), max_x=float(mxx), max_y=float(mxy)))
        dt = time.perf_counter() - t
        out["cpu_baseline"] = {"value": nc / dt, "unit": "frames/s", "cores": 1, "kind": "port",
                               "sample": f"{nc} of the same frames, oracle undistort + grid through ctypes"}
    ex.close()
    return out


def bench_bow(args, dev, local, matcher):
    """N2: Frame::ComputeBoW (src/Frame.cc:279-286) for a set of resident frames with an ORBvoc-sized tree
    (k = 10, L = 6: 1,111,111 nodes, 10^6 words; synthetic, the vocabulary file is not in the reference tree).
    CPU column: the reference's own DBoW2 transform (oracle/_ref, compiled from the reference sources) when that
    library travelled with the snapshot, else the pinned oracle port; one host thread, as in the reference."""
    import torch
    from oracle import bow_oracle as B
    from orb_slam_b200 import MatchFrame, ORBVocabulary
    F, NF = 512, 1000
    V = B.synth_vocabulary_fast(k=10, L=6, seed=1)
    keys = ("k", "L", "scoring", "weighting", "parent", "desc", "weight", "is_leaf")
    G = ORBVocabulary.from_arrays(matcher, **{k: V[k] for k in keys})
    rng = np.random.default_rng(5)
    leaves = V["desc"][V["is_leaf"] == 1]
    frames = []
    for f in range(F):
        d = leaves[rng.integers(0, len(leaves), NF)].copy()
        noise = rng.integers(0, 256, (NF, 3))
        for c in range(3):
            d[np.arange(NF), noise[:, c] >> 3] ^= (1 << (noise[:, c] & 7)).astype(np.uint8)
        z = np.zeros(NF, np.float32)
        frames.append(MatchFrame(z, z, z, np.zeros(NF, np.int32), d))
    fs = matcher.upload(frames)
    for _ in range(3):
        fs.compute_bow(G, 4)
    kms = []
    t = time.perf_counter()
    steps = max(args.steps, 5)
    for _ in range(steps):
        fs.compute_bow(G, 4)
        kms.append(matcher.last_kernel_ms())
    ms = (time.perf_counter() - t) * 1e3 / steps
    kms = float(np.mean(kms))
    out = {"metric": "ComputeBoW frames/s (1000 descriptors/frame, k=10 L=6 vocabulary, levelsup 4)",
           "value": F / (kms * 1e-3), "unit": "frames/s", "kernel_ms_per_step": kms, "frames_per_step": F,
           "e2e": {"value": F / (ms * 1e-3), "unit": "frames/s",
                   "note": "wall clock of orbm_frameset_compute_bow incl. its allocations and the offsets read-back"},
           "descriptor_distances_per_s": F * NF * 60 / (kms * 1e-3)}
    if not args.no_cpu:
        nc = 16
        if B.have_ref():
            import tempfile
            V2 = B.synth_vocabulary_fast(k=10, L=4, seed=1)     # text round trip of a 10^6-word file takes minutes
            with tempfile.TemporaryDirectory() as td:
                p = td + "/voc.txt"
                B.save_text(p, **V2)
                R = B.RefVocabulary.load_text(p)
            t = time.perf_counter()
            for f in range(nc):
                R.transform(frames[f].desc, 2)
            dt = time.perf_counter() - t
            out["cpu_baseline"] = {"value": nc / dt, "unit": "frames/s", "cores": 1, "kind": "reference",
                                   "sample": f"{nc} frames through the reference's DBoW2 transform (oracle/_ref) with a "
                                             "k=10 L=4 tree (40 distances/descriptor instead of 60: an upper bound "
                                             "on its L=6 rate)"}
        else:
            O = B.Vocabulary.from_arrays(**{k: V[k] for k in keys}, fast=True)
            t = time.perf_counter()
            for f in range(nc):
                O.transform(frames[f].desc, 4)
            dt = time.perf_counter() - t
            out["cpu_baseline"] = {"value": nc / dt, "unit": "frames/s", "cores": 1, "kind": "port",
                                   "sample": f"{nc} of the same frames, oracle transform, same tree"}
    fs.close(); G.close()
    return out


def bench_distinctive(args, dev, local, matcher):
    """N4: MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:185-250) for a batch of map points with
    2..30 observations each (what LocalMapping re-runs for every point a new keyframe touches)."""
    from oracle import match_oracle as MO
    rng = np.random.default_rng(17)
    P = 200_000
    sizes = rng.integers(2, 31, P)
    start = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    desc = rng.integers(0, 256, (int(start[-1]), 32), dtype=np.uint8)
    for _ in range(2):
        matcher.distinctive_descriptors(desc, start)
    kms = []
    t = time.perf_counter()
    steps = max(args.steps, 5)
    for _ in range(steps):
        best, _ = matcher.distinctive_descriptors(desc, start)
        kms.append(matcher.last_kernel_ms())
    ms = (time.perf_counter() - t) * 1e3 / steps
    kms = float(np.mean(kms))
    pairs = float((sizes.astype(np.int64) ** 2).sum())
    out = {"metric": "ComputeDistinctiveDescriptors map points/s (2..30 observations per point)",
           "value": P / (kms * 1e-3), "unit": "points/s", "kernel_ms_per_step": kms, "points_per_step": P,
           "descriptor_distances_per_s": pairs / (kms * 1e-3),
           "e2e": {"value": P / (ms * 1e-3), "unit": "points/s",
                   "h2d_bytes_per_step": int(desc.nbytes + start.nbytes), "d2h_bytes_per_step": P * 36}}
    if not args.no_cpu:
        nc = 20_000
        t = time.perf_counter()
        ref = MO.distinctive_descriptors(desc[: start[nc]], start[: nc + 1])
        dt = time.perf_counter() - t
        out["cpu_baseline"] = {"value": nc / dt, "unit": "points/s", "cores": 1, "kind": "port",
                               "sample": f"first {nc} of the same points, oracle (-O2), one thread"}
        out["gpu_equals_cpu"] = bool(np.array_equal(best[:nc], ref))
    return out


def bench_frustum(args, dev, local, matcher):
    """N3: Frame::isInFrustum (src/Frame.cc:136-197) of a 20,000-point local map against 64 views."""
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "tests"))
    import matchdata as MD
    from oracle import frame_oracle as FO
    from orb_slam_b200 import FrustumView
    NV, NP = 64, 20_000
    scenes = [MD.frustum_scene(s, n=NP) for s in range(NV)]
    pos, nrm, mn, mx = scenes[0][5:]
    oviews = [FO.make_view(Rcw, tcw, *K, b, sf) for Rcw, tcw, K, b, sf, *_ in scenes]
    views = [FrustumView.make(sc[0], sc[1], np.array(ov.Ow, np.float32), *sc[2], sc[3], sc[4])
             for sc, ov in zip(scenes, oviews)]
    for _ in range(2):
        matcher.is_in_frustum(views, pos, nrm, mn, mx)
    kms = []
    steps = max(args.steps, 5)
    t = time.perf_counter()
    for _ in range(steps):
        res = matcher.is_in_frustum(views, pos, nrm, mn, mx)
        kms.append(matcher.last_kernel_ms())
    ms = (time.perf_counter() - t) * 1e3 / steps
    kms = float(np.mean(kms))
    out = {"metric": "isInFrustum (view, map point) tests/s", "value": NV * NP / (kms * 1e-3), "unit": "tests/s",
           "kernel_ms_per_step": kms, "tests_per_step": NV * NP,
           "e2e": {"value": NV * NP / (ms * 1e-3), "unit": "tests/s", "h2d_bytes_per_step": NP * 32 + NV * 148,
                   "d2h_bytes_per_step": NV * NP * 17}}
    if not args.no_cpu:
        t = time.perf_counter()
        ref = [FO.is_in_frustum(ov, pos, nrm, mn, mx) for ov in oviews[:8]]
        dt = time.perf_counter() - t
        out["cpu_baseline"] = {"value": 8 * NP / dt, "unit": "tests/s", "cores": 1, "kind": "port",
                               "sample": "8 of the views, oracle (-O2), one thread"}
        out["gpu_equals_cpu"] = bool(all(np.array_equal(res[0][w], ref[w][0]) and
                                         np.array_equal(res[1][w].view(np.uint32), ref[w][1].view(np.uint32))
                                         for w in range(8)))
    return out


def bench_chain(args, dev, local, matcher):
    """The relocalisation / loop-detection front-end as ONE device-resident chain, per step 128 frame pairs (two shifted
    640x480 views of one scene each): ORB extraction (E0-E10) -> Frame glue (N1: undistort + grid) -> ComputeBoW (N2,
    k=10 L=6 vocabulary) -> SearchByBoW between the two views of every pair (M5, kf_kf, ratio 0.75, orientation
    check).  Images start in HBM; only the match rows and counts come back.  CPU column: the same chain through the
    oracle on one host thread (what one reference thread does), results compared."""
    import torch
    from oracle import bow_oracle as B
    from orb_slam_b200 import Camera, FrameSet, KP_DTYPE, ORBextractor, ORBmatcher, ORBVocabulary, synth_frames_device
    W, H, BW, BH, NF, NP = 640, 480, 704, 544, 1000, 128
    big = torch.empty((NP, BH, BW), dtype=torch.uint8, device=dev)
    synth_frames_device(big.data_ptr(), NP, BW, BH, BW, BH * BW, seed0=70_000)
    rng = np.random.default_rng(9)
    sh = rng.integers(-30, 31, (NP, 2))
    frames = torch.empty((2 * NP, H, W), dtype=torch.uint8, device=dev)
    for p in range(NP):
        frames[2 * p] = big[p, 32:32 + H, 32:32 + W]
        frames[2 * p + 1] = big[p, 32 + sh[p, 1]:32 + sh[p, 1] + H, 32 + sh[p, 0]:32 + sh[p, 0] + W]
    del big
    cam = (268.9633 * 2, 269.9858 * 2, 157.6087 * 2, 114.6369 * 2, -0.4157, 0.2624, 0.0, 0.0, -0.1178, W, H)
    camera = Camera(*cam)
    V = B.synth_vocabulary_fast(k=10, L=6, seed=1)
    keys = ("k", "L", "scoring", "weighting", "parent", "desc", "weight", "is_leaf")
    voc = ORBVocabulary.from_arrays(matcher, **{k: V[k] for k in keys})
    F = 2 * NP
    ex = ORBextractor(NF, 1.2, 8, 1, 20, max_width=W, max_height=H, max_batch=F, device=local)
    d_k = torch.zeros((F, NF, 28), dtype=torch.uint8, device=dev)
    d_d = torch.zeros((F, NF, 32), dtype=torch.uint8, device=dev)
    d_n = torch.zeros(F, dtype=torch.int32, device=dev)
    stream = torch.cuda.Stream(device=dev)
    om = ORBmatcher(0.75, True, matcher=matcher)
    i1 = np.arange(0, F, 2, dtype=np.int32); i2 = i1 + 1

    def step(keep=False):
        ex.extract_batch_device(frames.data_ptr(), F, W, H, W, H * W, d_k.data_ptr(), d_d.data_ptr(), d_n.data_ptr(),
                                stream.cuda_stream)
        fs = FrameSet.from_extractor(matcher, d_k.data_ptr(), d_d.data_ptr(), d_n.data_ptr(), F, NF, camera,
                                     stream.cuda_stream)
        fs.fill_valid(True)
        fs.compute_bow(voc, 4)
        nm, rows = om.SearchByBoW(fs, fs, i1, i2, "kf_kf")
        if keep:
            return fs, nm, rows
        fs.close()
        return None, nm, rows
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    steps = max(args.steps, 5)
    t = time.perf_counter()
    for _ in range(steps):
        _, nm, rows = step()
    ms = (time.perf_counter() - t) * 1e3 / steps
    out = {"metric": "front-end chain frames/s (extract -> frame glue -> ComputeBoW -> SearchByBoW, device resident)",
           "value": F / (ms * 1e-3), "unit": "frames/s", "ms_per_step": ms, "frames_per_step": F,
           "pairs_per_step": NP, "mean_matches_per_pair": float(nm.mean()), "d2h_bytes_per_step": int(rows.nbytes + nm.nbytes + 3 * 4 * F),
           "note": "wall clock incl. all allocations and syncs of the four calls; images resident in HBM"}
    if not args.no_cpu:
        from oracle import frame_oracle as FO
        from oracle import match_oracle as MO
        from oracle import oracle as O
        from orb_slam_b200 import MatchFrame
        nc = 4
        himg = frames[: 2 * nc].cpu().numpy()
        ovoc = B.Vocabulary.from_arrays(**{k: V[k] for k in keys}, fast=True)
        E = O.OracleExtractor(NF, 1.2, 8, 1, 20, fast_build=True)
        ocam = FO.Camera(*cam[:9])
        (mnx, mxx, mny, mxy), _ = FO.image_bounds(ocam, W, H)
        t = time.perf_counter()
        ok = True
        for p in range(nc):
            fr = []
            for f in (2 * p, 2 * p + 1):
                k, d = E(himg[f])
                un = FO.undistort_points(ocam, np.stack([k["x"], k["y"]], 1))
                _, fv = ovoc.transform(d, 4)
                fr.append(MatchFrame(un[:, 0], un[:, 1], k["angle"], k["octave"], d, np.ones(len(k), np.uint8),
                                     float(mnx), float(mny), float(mxx), float(mxy), fv))
            on, orow = MO.search_by_bow(fr[0], fr[1], "kf_kf", 0.75, True)
            ok = ok and on == nm[p] and np.array_equal(rows[p][: len(orow)], orow)
        dt = time.perf_counter() - t
        out["cpu_baseline"] = {"value": 2 * nc / dt, "unit": "frames/s", "cores": 1, "kind": "port",
                               "sample": f"{nc} of the same pairs through the oracle chain (-O3 -march=native)"}
        out["gpu_equals_cpu"] = bool(ok)
    ex.close(); voc.close()
    return out
