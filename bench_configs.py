"""Optional leg of bench.py (`python bench.py --all-configs`): BASELINE.json configs 1 and 3 (the bench.py line is
config 2; configs 4/5 are bench.py's "matching" block).
  config 1  one 640x480 frame through orbx_extract (host buffers): latency per call and frames/s
  config 3  1241x376 frames, 2000 features: batched throughput, device-resident and end-to-end
CPU columns: bench.py's cpu_baseline leg — the oracle port (oracle/, -O3 -march=native) on this box's host cores.
Returns a dict that bench.py embeds under "other_configs".
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def run(with_cpu=True):
    import torch
    from orb_slam_b200 import ORBextractor, synth_frames_device
    from oracle import oracle as O
    threads = os.cpu_count() or 1
    out = {}
    # ---- config 1 --------------------------------------------------------------------------------------------------
    img = O.synth(480, 640, 1)
    ex = ORBextractor(1000, 1.2, 8, 1, 20, max_width=640, max_height=480, max_batch=1)
    for _ in range(20):
        ex(img)
    t = time.perf_counter()
    n = 300
    for _ in range(n):
        k, d = ex(img)
    dt = (time.perf_counter() - t) / n
    E = O.OracleExtractor(1000, 1.2, 8, 1, 20, fast_build=True)
    E(img)
    t = time.perf_counter()
    for _ in range(20):
        ok, od = E(img)
    cdt = (time.perf_counter() - t) / 20
    out["config1_single_frame"] = {"gpu_ms_per_call": dt * 1e3, "gpu_frames_s": 1 / dt, "cpu_1thread_ms": cdt * 1e3,
                                   "cpu_1thread_frames_s": 1 / cdt, "keypoints": int(len(k)),
                                   "bit_exact": bool(np.array_equal(d, od) and len(k) == len(ok)),
                                   "note": "orbx_extract: pageable numpy in/out, H2D + 11 kernels + D2H + sync per call"}
    ex.close()
    # ---- config 3 --------------------------------------------------------------------------------------------------
    W, H, NF, F = 1241, 376, 2000, 2048
    ex = ORBextractor(NF, 1.2, 8, 1, 20, max_width=W, max_height=H, max_batch=256)
    pitch = 1248
    d_frames = torch.empty((F, H, pitch), dtype=torch.uint8, device="cuda")
    synth_frames_device(d_frames.data_ptr(), F, W, H, pitch, H * pitch, seed0=7000)
    d_k = torch.empty((F, NF, 28), dtype=torch.uint8, device="cuda")
    d_d = torch.empty((F, NF, 32), dtype=torch.uint8, device="cuda")
    d_n = torch.zeros(F, dtype=torch.int32, device="cuda")
    stream = torch.cuda.Stream()

    def step():
        ex.extract_batch_device(d_frames.data_ptr(), F, W, H, pitch, H * pitch, d_k.data_ptr(), d_d.data_ptr(),
                                d_n.data_ptr(), stream.cuda_stream)
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(5):
        step()
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    h_frames = torch.empty((F, H, W), dtype=torch.uint8, pin_memory=True)   # tight rows, like a cv::Mat
    h_frames.copy_(d_frames[:, :, :W])
    h_k = torch.empty((F, NF, 28), dtype=torch.uint8, pin_memory=True)
    h_d = torch.empty((F, NF, 32), dtype=torch.uint8, pin_memory=True)
    h_n = torch.zeros(F, dtype=torch.int32, pin_memory=True)

    def step_e2e():
        ex.extract_batch_ptr(h_frames.data_ptr(), F, W, H, W, H * W, h_k.data_ptr(), h_d.data_ptr(), h_n.data_ptr())
    step_e2e()
    t = time.perf_counter()
    for _ in range(3):
        step_e2e()
    torch.cuda.synchronize()
    e2e = (time.perf_counter() - t) / 3
    frames = np.ascontiguousarray(h_frames.numpy()[:8 * threads])
    E = O.OracleExtractor(NF, 1.2, 8, 1, 20, fast_build=True)
    E.extract_many(frames[:threads], threads)
    t = time.perf_counter()
    _, cd, cn = E.extract_many(frames, threads)
    cdt = time.perf_counter() - t
    t = time.perf_counter()
    E.extract_many(frames[:8], 1)
    c1 = time.perf_counter() - t
    nn = h_n.numpy()
    same = all(np.array_equal(h_d.numpy()[f, :nn[f]], cd[f, :cn[f]]) for f in range(len(frames)))
    out["config3_1241x376_2000kp"] = {"gpu_device_frames_s": F / (ms * 1e-3), "gpu_e2e_frames_s": F / e2e,
                                      "cpu_all_threads_frames_s": len(frames) / cdt, "cpu_threads": threads,
                                      "cpu_1thread_frames_s": 8 / c1, "keypoints_per_frame": float(nn.mean()),
                                      "descriptors_bit_exact_on_cpu_sample": bool(same)}
    return out
