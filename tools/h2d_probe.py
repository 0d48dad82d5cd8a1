"""Pinned host <-> device copy bandwidth of this box (the ceiling of the e2e leg): python tools/h2d_probe.py"""
import torch, time
n = 3 * 1024**3
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
h2 = torch.empty(600 * 1024**2, dtype=torch.uint8).pin_memory()
d2 = torch.empty(600 * 1024**2, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def t(fn, reps=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps
a = t(lambda: d.copy_(h, non_blocking=True))
b = t(lambda: h2.copy_(d2, non_blocking=True))
def both():
    with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2):
        for _ in range(5): h2.copy_(d2, non_blocking=True)
c = t(both)
print(f"H2D {n/a/1e9:.1f} GB/s  D2H {h2.numel()/b/1e9:.1f} GB/s  duplex: H2D {n/c/1e9:.1f} GB/s + D2H {5*h2.numel()/c/1e9:.1f} GB/s")
