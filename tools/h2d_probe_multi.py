"""Aggregate pinned host -> device bandwidth with N GPUs copying at once (the ceiling of the e2e leg at N > 1).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/h2d_probe_multi.py

Every rank copies the byte counts one bench step moves (3.07 GB in, 0.6 GB out, both directions in flight) between its
own pinned buffers and its own GPU; all ranks start together and rank 0 prints the sum over ranks.  With ORBX_NUMA=1 each
rank first binds itself to the CPUs nvidia-smi lists as its GPU's affinity (pinned pages then come from that node).
"""
import os, subprocess, time
import torch
import torch.distributed as dist

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
bound = ""
if os.environ.get("ORBX_NUMA") == "1":
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64 + 8)
        cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1}
        ok = cpus & os.sched_getaffinity(0)
        if ok:
            os.sched_setaffinity(0, ok)
            bound = f"{len(ok)} cpus"
        else:
            bound = "gpu's cpus outside this cpuset"
    except Exception as e:  # report, never hide
        bound = f"failed: {e!r}"
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n_in, n_out = 3_072_000_000, 600_040_000
h_in, d_in = torch.empty(n_in, dtype=torch.uint8).pin_memory(), torch.empty(n_in, dtype=torch.uint8, device="cuda")
h_out, d_out = torch.empty(n_out, dtype=torch.uint8).pin_memory(), torch.empty(n_out, dtype=torch.uint8, device="cuda")
h_in.fill_(1); h_out.fill_(1)                     # touch the pages from this rank's CPUs
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def step(duplex):
    with torch.cuda.stream(s1):
        d_in.copy_(h_in, non_blocking=True)
    if duplex:
        with torch.cuda.stream(s2):
            h_out.copy_(d_out, non_blocking=True)


def timed(duplex, reps=5):
    step(duplex); torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        step(duplex)
    torch.cuda.synchronize()
    t = torch.tensor([(time.perf_counter() - t0) / reps], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


a, b = timed(False), timed(True)
info = [None] * world
dist.all_gather_object(info, (local, bound))
if rank == 0:
    print(f"N={world} numa={os.environ.get('ORBX_NUMA', '0')} H2D alone {world * n_in / a / 1e9:.1f} GB/s aggregate; duplex: "
          f"H2D {world * n_in / b / 1e9:.1f} + D2H {world * n_out / b / 1e9:.1f} GB/s aggregate "
          f"(= {world * 10000 / b:.0f} frames/s if the copies were the whole step)  bind={info}")
dist.destroy_process_group()
