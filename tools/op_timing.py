"""Warm per-call timings of the small building blocks (development tool)."""
import sys
import torch
sys.path.insert(0, ".")
from linalg_b200 import dev  # noqa: E402


def t(fn, reps=20):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3  # us


torch.manual_seed(0)
info = torch.zeros(1, dtype=torch.int32, device="cuda")
for m in (512, 2048, 4096, 8192, 16384):
    for jb in (32,):
        A = torch.rand((jb, m), dtype=torch.float64, device="cuda")
        ipiv = torch.zeros(jb, dtype=torch.int32, device="cuda")
        print(f"lu panel m={m} jb={jb}: {t(lambda: dev.getrf(m, jb, A, m, ipiv, info)):.1f} us", flush=True)
for m in (512, 4096, 16384):
    for na in (32, 64, 128):
        T = torch.rand((na, na), dtype=torch.float64, device="cuda") + na * torch.eye(na, dtype=torch.float64, device="cuda")
        B = torch.rand((na, m), dtype=torch.float64, device="cuda")   # m x na col-major
        print(f"trsm Right/Lower/C m={m} na={na}: {t(lambda: dev.trsm(dev.RIGHT, dev.LOWER, dev.CONJ_TRANS, dev.NON_UNIT, m, na, 1.0, T, na, B, m)):.1f} us", flush=True)
        B2 = torch.rand((m, na), dtype=torch.float64, device="cuda")  # na x m col-major
        print(f"trsm Left/Lower/N/Unit n={m} na={na}: {t(lambda: dev.trsm(dev.LEFT, dev.LOWER, dev.NO_TRANS, dev.UNIT, na, m, 1.0, T, na, B2, na)):.1f} us", flush=True)
for n in (32, 64, 128):
    G = torch.rand((n, n), dtype=torch.float64, device="cuda")
    S = G @ G.T + n * torch.eye(n, dtype=torch.float64, device="cuda")
    W = S.clone()
    def f():
        W.copy_(S)
        dev.potrf(dev.LOWER, n, W, n, info)
    tc = t(lambda: W.copy_(S))
    print(f"potrf leaf n={n}: {t(f) - tc:.1f} us (copy {tc:.1f})", flush=True)
for (m, n, k) in [(16384, 32, 32), (16384, 64, 64), (16384, 128, 128), (16384, 256, 256), (8192, 512, 128), (15360, 15360, 512), (128, 128, 128), (256, 256, 64)]:
    A = torch.rand((k, m), dtype=torch.float64, device="cuda")
    B = torch.rand((n, k), dtype=torch.float64, device="cuda")
    Cm = torch.zeros((n, m), dtype=torch.float64, device="cuda")
    us = t(lambda: dev.gemm(dev.NO_TRANS, dev.NO_TRANS, m, n, k, -1.0, A, m, B, k, 1.0, Cm, m), reps=10)
    print(f"gemm NN {m}x{n}x{k}: {us:.1f} us  {2*m*n*k/us*1e-6:.2f} TFLOP/s", flush=True)
