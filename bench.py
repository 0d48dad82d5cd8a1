#!/usr/bin/env python
"""bench.py — ORB extraction throughput on B200 (BASELINE.json metric, configs[1]).

  python bench.py --gpus N --steps K --warmup W          # N>1: launched by torchrun, one rank per GPU
  python bench.py --impl reference ...                   # the reference's CPU path (oracle restatement) on host cores

A step = one pass of the extractor over the rank's shard of the synthetic 640x480 sequence (10,000 frames per GPU,
weak scaling; frames are independent, so there is no data-path collective).  `value` = frames/s with the frames
already resident in HBM (CUDA events on the launching stream, max over ranks); `e2e` = the same pass through the
C-ABI call a user makes (orbx_extract_batch) with HOST pinned buffers, H2D of the frames and D2H of keypoints +
descriptors inside the timed region.  Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H, NFEAT = 640, 480, 1000
# algorithmic bytes per 640x480 frame, every stage reading its input and writing its output once
ALGO = {"pyramid": 926_546 + 643_332, "fast": 950_532, "select": 0, "blur": 1_901_064,
        "describe": 749_000 + 512_000 + 32_000 + 28_000}
ALGO_FRAME = 5_742_474
assert sum(ALGO.values()) == ALGO_FRAME
METRIC = "ORB extract frames/s (640x480, 1000 kp)"


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    def __init__(self, index):
        self.rows = []
        self.proc = None
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, line in self.rows:
            if t < t0 or t > t1 + 0.2:
                continue
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except Exception:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def _cpu_arms(frames, threads, reps=1, warm=True):
    """The reference's CPU path on `threads` host threads over `frames` [F,H,W], two ways: (i) "reference" — the
    reference's own src/ORBextractor.cc compiled in place (oracle/_ref/libref_orbextractor_native.so, its -O3) with its
    five OpenCV routines served by the oracle's cv2-pinned scalar restatements; (ii) "port" — the oracle restatement
    built with the reference's flags (-O3 -march=native).  Returns {kind: (frames/s, mean keypoints)}.  Real OpenCV is
    SIMD-optimised, the bridge's primitives are not, so the compiled reference here is SLOWER than it would be with
    OpenCV; the faster arm is the conservative baseline and is the one reported as `value`."""
    from oracle import oracle as O
    from oracle import ref_extractor as RE
    out = {}
    arms = [("port", O.OracleExtractor(NFEAT, 1.2, 8, 1, 20, fast_build=True))]
    if RE.native_available():
        arms.append(("reference", RE.NativeRefExtractor(NFEAT, 1.2, 8, 1, 20)))
    for kind, E in arms:
        if warm:
            E.extract_many(frames[:min(threads, len(frames))], threads)
        t = time.perf_counter()
        for _ in range(reps):
            _, _, n = E.extract_many(frames, threads)
        out[kind] = (len(frames) * reps / (time.perf_counter() - t), float(n.mean()))
    return out


def _cpu_pick(arms):
    kind = max(arms, key=lambda k: arms[k][0])
    extra = {f"{k}_value": v[0] for k, v in arms.items()}
    return kind, arms[kind][0], extra


def cpu_baseline(nframes, threads, d_frames_cpu):
    """CPU baseline on the box's host cores, bounded sample (see _cpu_arms)."""
    arms = _cpu_arms(d_frames_cpu[:nframes], threads)
    kind, v, extra = _cpu_pick(arms)
    n1 = max(nframes // (2 * threads), 8)
    one = _cpu_arms(d_frames_cpu[:n1], 1, warm=False)
    return {"value": v, "unit": "frames/s", "cores": threads, "kind": kind,
            "sample": f"{nframes} of the workload's 640x480 frames, {threads} host threads; both CPU arms timed, the "
                      f"faster one reported (reference = src/ORBextractor.cc compiled in place, OpenCV routines from "
                      f"the oracle's scalar restatements; port = oracle/liborb_oracle_fast.so, -O3 -march=native)",
            **extra, "single_thread_value": one[kind][0], "keypoints_per_frame": arms[kind][1]}


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the path on all host threads, bounded sample per
    step.  Both CPU arms of _cpu_arms are timed; the faster is the line's value (cpu_baseline.kind says which)."""
    if rank != 0:
        return
    import numpy as np
    from oracle import oracle as O
    threads = os.cpu_count() or 1
    nframes = args.ref_frames
    frames = np.stack([O.synth(H, W, s) for s in range(nframes)])
    for _ in range(args.warmup):
        _cpu_arms(frames[:threads], threads, warm=False)
    arms = _cpu_arms(frames, threads, reps=args.steps, warm=False)
    kind, v, extra = _cpu_pick(arms)
    dt = nframes / v
    sample = (f"{nframes} frames per step of the 640x480 synthetic sequence, {threads} host threads; both CPU arms timed "
              f"(reference = src/ORBextractor.cc compiled in place with scalar OpenCV stand-in routines, port = oracle "
              f"restatement with the reference's flags), the faster one reported")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "configs[1]: batched ORB extraction, synthetic 640x480 sequence, 1000 features, 8 levels, "
                               "scale 1.2, fastTh 20 (bounded CPU sample)", "frames_per_step": nframes},
        "cpu_baseline": {"value": v, "unit": "frames/s", "cores": threads, "kind": kind, "sample": sample, **extra},
        "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--frames", type=int, default=10_000, help="frames per GPU per step")
    ap.add_argument("--batch", type=int, default=256, help="frames per kernel wave (orbx max_batch)")
    ap.add_argument("--cpu-frames", type=int, default=0, help="cpu_baseline sample (0 = auto ~15 s)")
    ap.add_argument("--ref-frames", type=int, default=512)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-matching", action="store_true", help="skip the Hamming-matching legs")
    ap.add_argument("--query-keyframes", type=int, default=8, help="config 5: query keyframes (x1000 descriptors) per step")
    ap.add_argument("--match-pairs", type=int, default=2048, help="config 4: frame pairs per GPU per step")
    ap.add_argument("--host-wc", action="store_true", help="e2e leg: input frames in WRITE-COMBINED pinned memory (orbx_host_alloc)")
    ap.add_argument("--no-other-configs", action="store_true", help="skip BASELINE configs 1 and 3 (rank 0, N=1 only)")
    ap.add_argument("--all-configs", action="store_true", help=argparse.SUPPRESS)   # kept for old command lines
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import numpy as np
    import torch
    import torch.distributed as dist
    from orb_slam_b200 import ORBextractor, synth_frames_device

    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    F = args.frames
    ex = ORBextractor(NFEAT, 1.2, 8, 1, 20, max_width=W, max_height=H, max_batch=args.batch, device=local)

    # inputs: synthetic frames generated on the device (seed = global frame index), resident in HBM
    d_frames = torch.empty((F, H, W), dtype=torch.uint8, device=dev)
    synth_frames_device(d_frames.data_ptr(), F, W, H, W, H * W, seed0=rank * F)
    d_kps = torch.empty((F, NFEAT, 28), dtype=torch.uint8, device=dev)
    d_desc = torch.empty((F, NFEAT, 32), dtype=torch.uint8, device=dev)
    d_n = torch.zeros(F, dtype=torch.int32, device=dev)
    # host side of the e2e leg: pinned frames in, pinned results out
    wc_ptr = None
    if args.host_wc:
        import ctypes
        from orb_slam_b200._lib import lib as _orbx_lib
        wc_ptr = ctypes.c_void_p()
        assert _orbx_lib().orbx_host_alloc(F * H * W, 1, ctypes.byref(wc_ptr)) == 0
        h_frames = torch.from_numpy(np.ctypeslib.as_array(ctypes.cast(wc_ptr, ctypes.POINTER(ctypes.c_uint8)), shape=(F, H, W)))
        stage = torch.empty((min(F, 512), H, W), dtype=torch.uint8, pin_memory=True)
        for f0 in range(0, F, stage.shape[0]):                        # fill through a pinned bounce buffer (CPU WRITES only)
            n_ = min(stage.shape[0], F - f0)
            stage[:n_].copy_(d_frames[f0:f0 + n_])
            h_frames[f0:f0 + n_] = stage[:n_]
        del stage
    else:
        h_frames = torch.empty((F, H, W), dtype=torch.uint8, pin_memory=True)
        h_frames.copy_(d_frames)
    h_kps = torch.empty((F, NFEAT, 28), dtype=torch.uint8, pin_memory=True)
    h_desc = torch.empty((F, NFEAT, 32), dtype=torch.uint8, pin_memory=True)
    h_n = torch.zeros(F, dtype=torch.int32, pin_memory=True)
    # a real (non-default) stream: the library treats a NULL stream as "use the handle's own stream", and CUDA
    # events must be recorded on the stream the kernels are launched on
    stream = torch.cuda.Stream(device=dev)

    def step_device():
        ex.extract_batch_device(d_frames.data_ptr(), F, W, H, W, H * W, d_kps.data_ptr(), d_desc.data_ptr(),
                                d_n.data_ptr(), stream.cuda_stream)

    def step_e2e():
        ex.extract_batch_ptr(h_frames.data_ptr(), F, W, H, W, H * W, h_kps.data_ptr(), h_desc.data_ptr(),
                             h_n.data_ptr())

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def maxr(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-resident leg ------------------------------------------------------------------------------------
    for _ in range(args.warmup):
        step_device()
    ex.synchronize()
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    l0 = ex.kernel_launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    e0.record(stream)
    for _ in range(args.steps):
        step_device()
    e1.record(stream)
    barrier()
    t1 = time.time()
    launches = ex.kernel_launches() - l0
    ex.synchronize()
    ms_step = maxr(e0.elapsed_time(e1) / args.steps)
    clocks = sampler.stop(t0, t1) if sampler else None
    kp_mean = float(d_n.float().mean().item())

    # ---- per-stage CUDA-event timing of one pass (dominant kernel for the roofline) ------------------------------
    ex.profile_enable(True)
    ex.profile_read()
    step_device()
    ex.synchronize()
    prof = ex.profile_read()
    ex.profile_enable(False)
    peak, peak_src = peaks()
    dom = max(prof, key=lambda s: prof[s][0])
    dom_ms, dom_launches = prof[dom]
    dom_gbs = ALGO[dom] * F / (dom_ms * 1e-3) / 1e9
    # per-kernel counters of one ncu --set full capture (profiles/r2_ncu_full_extract.csv -> tools/ncu_counters.py):
    # executed warp instructions and DRAM bytes per frame do not depend on clocks or co-running kernels, so together with
    # the LIVE stage times they give the roof these integer kernels actually sit under: the SM's issue rate (4 warp
    # instructions / clk / SM) and the busiest execution pipe (ALU class: 2 / clk / SM, tools/int_ops_microbench.cu)
    counters = {}
    try:
        counters = json.load(open(os.path.join(ROOT, "profiles", "r2_kernel_counters.json")))["stages"]
    except Exception:
        pass
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    clk = ((clocks or {}).get("sm_mhz") or 1965.0) * 1e6
    stages = {}
    for s_ in prof:
        t = prof[s_][0] * 1e-3
        e = {"ms": round(prof[s_][0], 3)}
        if ALGO[s_]:
            e["hbm_frac"] = round(ALGO[s_] * F / t / 1e9 / peak, 4)
        c = counters.get(s_)
        if c and t > 0:
            e["issue_frac"] = round(c["warp_inst_per_frame"] * F / (sms * 4 * clk * t), 3)
            e["pipe"] = c["busiest_pipe"]
            e["pipe_frac"] = round(c["ncu_busiest_pipe_pct"] / 100 * (c["ncu_us_per_frame"] * 1e-6 * F) / t, 3)
            e["thread_inst_per_algo_byte"] = round(32 * c["warp_inst_per_frame"] / ALGO[s_], 1) if ALGO[s_] else None
            e["dram_over_algo"] = round(c["dram_bytes_per_frame"] / ALGO[s_], 2) if ALGO[s_] else None
        stages[s_] = e
    traffic = int(counters[dom]["dram_bytes_per_frame"] * args.batch) if dom in counters else None
    roofline = {"bound": "hbm", "kernel": dom, "achieved": dom_gbs, "peak": peak, "unit": "GB/s",
                "frac": dom_gbs / peak, "traffic": traffic, "peak_source": peak_src,
                "avg_launch_ms": dom_ms / max(dom_launches, 1), "launches_per_pass": dom_launches,
                "algorithmic_bytes_per_frame": ALGO[dom],
                "note": "frac is the contractual HBM fraction of the dominant kernel; the kernels are integer-issue bound "
                        "(stages[*].issue_frac = executed warp instructions / (SMs x 4 x clk x time), pipe_frac = busiest "
                        "execution pipe), traffic = DRAM bytes of one launch from ncu (profiles/r2_ncu_full_extract.csv)",
                "stage_ms_per_pass": {s: round(prof[s][0], 3) for s in prof}, "stages": stages,
                "whole_pipeline_frac": ALGO_FRAME * F / (ms_step * 1e-3) / 1e9 / peak}

    # ---- end-to-end leg (host pinned -> H2D -> kernels -> D2H) ----------------------------------------------------
    for _ in range(max(1, min(args.warmup, 2))):
        step_e2e()
    barrier()
    te = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    torch.cuda.synchronize()
    e2e_ms = maxr((time.perf_counter() - te) * 1e3 / args.steps)
    barrier()
    assert int(h_n.sum().item()) == int(d_n.sum().item()), "e2e and device-resident legs disagree"
    # ---- what the host <-> device fabric gives when the copies ARE the step: the same bytes, plain cudaMemcpyAsync, all
    # ranks at once, both directions in flight (the ceiling of the e2e leg; on a multi-GPU VM it is shared by the ranks)
    fabric = None
    if not args.host_wc:
        s_in, s_out = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
        d_out = torch.empty(F * (NFEAT * 60 + 4), dtype=torch.uint8, device=dev)
        h_out = torch.empty(F * (NFEAT * 60 + 4), dtype=torch.uint8, pin_memory=True)

        def copies():
            with torch.cuda.stream(s_in):
                d_frames.copy_(h_frames, non_blocking=True)
            with torch.cuda.stream(s_out):
                h_out.copy_(d_out, non_blocking=True)
        copies()
        barrier()
        tf = time.perf_counter()
        for _ in range(3):
            copies()
        torch.cuda.synchronize()
        fab_ms = maxr((time.perf_counter() - tf) * 1e3 / 3)
        barrier()
        fabric = {"h2d_gbs": world * F * H * W / fab_ms / 1e6, "d2h_gbs": world * F * (NFEAT * 60 + 4) / fab_ms / 1e6,
                  "ms_per_step": fab_ms, "frames_s_if_copies_were_the_step": world * F / (fab_ms * 1e-3)}
        del d_out, h_out

    # ---- strong-scaling leg: config 2 as BASELINE.json words it - ONE 10,000-frame sequence split 10,000/G per rank ----
    strong = None
    if world > 1:
        Fs = (F * (rank + 1)) // world - (F * rank) // world

        def step_strong():
            ex.extract_batch_device(d_frames.data_ptr(), Fs, W, H, W, H * W, d_kps.data_ptr(), d_desc.data_ptr(),
                                    d_n.data_ptr(), stream.cuda_stream)

        def step_strong_e2e():
            ex.extract_batch_ptr(h_frames.data_ptr(), Fs, W, H, W, H * W, h_kps.data_ptr(), h_desc.data_ptr(),
                                 h_n.data_ptr())
        for _ in range(args.warmup):
            step_strong()
        ex.synchronize()
        barrier()
        e0.record(stream)
        for _ in range(args.steps):
            step_strong()
        e1.record(stream)
        barrier()
        ex.synchronize()
        s_ms = maxr(e0.elapsed_time(e1) / args.steps)
        step_strong_e2e()
        barrier()
        te = time.perf_counter()
        for _ in range(args.steps):
            step_strong_e2e()
        torch.cuda.synchronize()
        s_e2e = maxr((time.perf_counter() - te) * 1e3 / args.steps)
        barrier()
        strong = {"frames_total": F, "frames_per_gpu": Fs, "value": F / (s_ms * 1e-3), "unit": "frames/s",
                  "ms_per_step": s_ms, "e2e_value": F / (s_e2e * 1e-3), "e2e_ms_per_step": s_e2e, "scaling": "strong"}

    cpu = None
    if rank == 0 and not args.no_cpu:
        threads = os.cpu_count() or 1
        ncpu = args.cpu_frames or max(64, min(F, 8 * threads * 4))
        cpu = cpu_baseline(min(ncpu, F), threads, h_frames.numpy())
    matching = None
    if not args.no_matching:
        import bench_matching as BM
        from orb_slam_b200 import Matcher
        del d_kps, d_desc, h_kps, h_desc
        mt = Matcher(local)
        s2 = ClockSampler(local) if rank == 0 else None
        tk0 = time.time()
        knn = BM.bench_knn2(args, rank, world, dev, mt, None)
        tk1 = time.time()
        kclk = s2.stop(tk0, tk1) if s2 else None
        if kclk and kclk.get("sm_mhz"):
            sms = torch.cuda.get_device_properties(dev).multi_processor_count
            peak = sms * BM.POPC_PER_CLK_PER_SM * kclk["sm_mhz"] * 1e6 * world
            knn["roofline"]["popc_bound"].update({"peak": peak / 1e12, "frac": knn["value"] * 8 / peak, "clock_mhz_assumed": kclk["sm_mhz"]})
            knn["clocks"] = kclk
        init = BM.bench_init_search(args, rank, world, dev, local, mt)
        if rank == 0 and not args.no_cpu:
            knn["cpu_baseline"] = BM.bench_knn2_cpu(args)
        matching = {"knn2_all_pairs": knn, "search_for_initialization": init}
        if rank == 0:
            matching["frame_glue"] = BM.bench_frame_glue(args, dev, local, mt)
            matching["bow"] = BM.bench_bow(args, dev, local, mt)
            matching["distinctive_descriptors"] = BM.bench_distinctive(args, dev, local, mt)
            matching["is_in_frustum"] = BM.bench_frustum(args, dev, local, mt)
            matching["frontend_chain"] = BM.bench_chain(args, dev, local, mt)
    other = None
    if not args.no_other_configs and world == 1:
        import bench_configs
        del d_frames, h_frames
        ex.close()
        other = bench_configs.run()
    if world > 1:
        dist.barrier()
    if rank == 0:
        total = F * world
        line = ({
            "metric": METRIC, "value": total / (ms_step * 1e-3), "unit": "frames/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "configs[1]: batched ORB extraction over a 10k-frame synthetic 640x480 sequence "
                                   "per GPU, Data/Settings.yaml knobs (1000 features, 8 levels, scale 1.2, fastTh 20, "
                                   "FAST score)", "frames_per_gpu": F, "frames_total": total, "batch": args.batch,
                       "parallelism": f"frames sharded over {world} GPU(s), no collective",
                       "l2": f"inputs ({F * H * W / 1e6:.0f} MB per GPU) exceed the 126 MB L2; no flush needed",
                       "keypoints_per_frame": kp_mean},
            "e2e": {"value": total / (e2e_ms * 1e-3), "unit": "frames/s", "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": F * H * W, "d2h_bytes_per_step": F * (NFEAT * 60 + 4),
                    "api": "orbx_extract_batch (C ABI, pinned host buffers), wall clock between device syncs",
                    "host_memory": "write-combined pinned (orbx_host_alloc)" if args.host_wc else "pinned",
                    "host_fabric": fabric,
                    "frac_of_host_fabric": (fab_ms / e2e_ms) if fabric else None},
            "gpu_launches": launches, "roofline": roofline, "cpu_baseline": cpu, "clocks": clocks,
            "strong_scaling": strong, "matching": matching, "other_configs": other})
        # compact digest, printed LAST so that it survives a truncated stdout tail: every number BASELINE.md quotes
        sm = {"device_fps": round(line["value"]), "e2e_fps": round(line["e2e"]["value"]), "n_gpus": world,
              "stage_ms": roofline["stage_ms_per_pass"], "roofline_frac": round(roofline["frac"], 4),
              "e2e_frac_of_host_fabric": round(line["e2e"]["frac_of_host_fabric"], 3) if line["e2e"]["frac_of_host_fabric"] else None,
              "issue_frac": {s_: v.get("issue_frac") for s_, v in roofline["stages"].items()},
              "pipe_frac": {s_: v.get("pipe_frac") for s_, v in roofline["stages"].items()}}
        if strong:
            sm.update({"strong_scaling_fps": round(strong["value"]), "strong_scaling_e2e_fps": round(strong["e2e_value"])})
        if matching:
            k, i = matching["knn2_all_pairs"], matching["search_for_initialization"]
            sm.update({"knn2_pairs_s": float(f"{k['value']:.4g}"), "knn2_ms": round(k["ms_per_step"], 3),
                       "knn2_frac_tensor_fp8": round(k["roofline"]["frac"], 3), "knn2_x_popc_bound": round(k["roofline"]["popc_bound"]["frac"], 2), "knn2_nranks": world,
                       "sharded_equals_single": (k.get("check") or {}).get("sharded_equals_single"),
                       "knn2_accepted_frac": (k.get("check") or {}).get("accepted_frac"),
                       "init_pairs_s": round(i["value"]), "init_e2e_pairs_s": round(i["e2e"]["value"]),
                       "init_gpu_equals_cpu": (i.get("cpu_baseline") or {}).get("gpu_equals_cpu")})
            for name in ("frame_glue", "bow", "distinctive_descriptors", "is_in_frustum", "frontend_chain"):
                if name in matching:
                    sm[name + "_per_s"] = float(f"{matching[name]['value']:.4g}")
        if other:
            c1, c3 = other["config1_single_frame"], other["config3_1241x376_2000kp"]
            sm.update({"config1_ms_per_call": round(c1["gpu_ms_per_call"], 4), "config1_bit_exact": c1["bit_exact"],
                       "config3_device_fps": round(c3["gpu_device_frames_s"]), "config3_e2e_fps": round(c3["gpu_e2e_frames_s"]),
                       "config3_bit_exact": c3["descriptors_bit_exact_on_cpu_sample"]})
        line["summary"] = sm
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
