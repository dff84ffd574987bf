"""Device-resident timings of the QR paths: dgels / zgels 2^20 x 512, dgeqrf tall and square."""
import sys
import torch
sys.path.insert(0, ".")
from linalg_b200 import dev


def timed(fn, setup, reps=3):
    best = 1e30
    for _ in range(reps):
        setup()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


info = torch.zeros(1, dtype=torch.int32, device="cuda")
for dt, let, fl in ((torch.float64, "d", 1), (torch.complex128, "z", 4)):
    m, n = 1 << 20, 512
    A0 = torch.rand((n, m), dtype=dt, device="cuda")
    b0 = torch.rand((1, m), dtype=dt, device="cuda")
    A, b = torch.empty_like(A0), torch.empty_like(b0)
    ms = timed(lambda: dev.gels(dev.NO_TRANS, m, n, 1, A, m, b, m, info), lambda: (A.copy_(A0), b.copy_(b0)))
    r = A0.T @ b[0, :n] - b0[0]
    res = float(((A0.conj() @ r).norm() / (A0.norm() * r.norm())).item())
    print(f"{let}gels 2^20x512: {ms:.2f} ms  {fl * (2 * m * n * n - 2 / 3 * n ** 3) / ms * 1e-9:.2f} TFLOP/s  normal-eq residual {res:.2e}", flush=True)
    tau = torch.zeros(n, dtype=dt, device="cuda")
    ms = timed(lambda: dev.geqrf(m, n, A, m, tau), lambda: A.copy_(A0))
    print(f"{let}geqrf 2^20x512: {ms:.2f} ms", flush=True)
    del A0, b0, A, b
    torch.cuda.empty_cache()
for n in (4096, 8192, 16384):
    A0 = torch.rand((n, n), dtype=torch.float64, device="cuda")
    A = torch.empty_like(A0)
    tau = torch.zeros(n, dtype=torch.float64, device="cuda")
    ms = timed(lambda: dev.geqrf(n, n, A, n, tau), lambda: A.copy_(A0), reps=2)
    print(f"dgeqrf {n}^2: {ms:.2f} ms  {4 / 3 * n ** 3 / ms * 1e-9:.2f} TFLOP/s", flush=True)
