"""Pinned host <-> device copy bandwidth on this box (development tool)."""
import time
import torch
n = 1 << 28          # 2 GiB of float64
h = torch.empty(n, dtype=torch.float64).pin_memory()
d = torch.empty(n, dtype=torch.float64, device="cuda")
h2 = torch.empty(n, dtype=torch.float64).pin_memory()
d2 = torch.empty(n, dtype=torch.float64, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
for name, fn in (("H2D", lambda: d.copy_(h, non_blocking=True)), ("D2H", lambda: h.copy_(d, non_blocking=True))):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    print(name, f"{3 * n * 8 / (time.perf_counter() - t0) / 1e9:.1f} GB/s")
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(3):
    with torch.cuda.stream(s1):
        d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2):
        h2.copy_(d2, non_blocking=True)
torch.cuda.synchronize()
dt = time.perf_counter() - t0
print(f"both directions at once: {3 * n * 8 / dt / 1e9:.1f} GB/s each way")
