"""Warm timings of one m x NB LU panel and its pieces (development tool)."""
import sys
import torch
sys.path.insert(0, ".")
from linalg_b200 import dev  # noqa: E402


def t(fn, reps=5):
    fn(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best * 1e3


torch.manual_seed(0)
info = torch.zeros(1, dtype=torch.int32, device="cuda")
for m in (16384, 8192, 2048):
    for nb in (32, 64, 128, 256, 512):
        A0 = torch.rand((nb, m), dtype=torch.float64, device="cuda")
        A = A0.clone()
        ipiv = torch.zeros(nb, dtype=torch.int32, device="cuda")
        def f():
            A.copy_(A0)
            dev.getrf(m, nb, A, m, ipiv, info)
        tc = t(lambda: A.copy_(A0))
        print(f"getrf panel m={m} nb={nb}: {t(f) - tc:.0f} us", flush=True)
