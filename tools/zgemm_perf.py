import sys, torch
sys.path.insert(0, ".")
from linalg_b200 import dev
def timeit(fn, reps=3):
    fn(); torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best
for n in (2048, 8192):
    A = torch.rand((n, n), dtype=torch.complex128, device="cuda")
    B = torch.rand((n, n), dtype=torch.complex128, device="cuda")
    C = torch.zeros((n, n), dtype=torch.complex128, device="cuda")
    ms = timeit(lambda: dev.gemm(dev.NO_TRANS, dev.NO_TRANS, n, n, n, 1.0, A, n, B, n, 0.0, C, n))
    ref = (A.T @ B.T).T if n <= 2048 else None
    err = ((C - ref).norm() / ref.norm()).item() if ref is not None else float("nan")
    print(f"zgemm n={n}: {ms:.2f} ms {8*n**3/ms*1e-9:.2f} TFLOP/s path={dev.last_gemm_path()} relerr={err:.2e}", flush=True)
    ipiv = torch.zeros(n, dtype=torch.int32, device="cuda"); info = torch.zeros(1, dtype=torch.int32, device="cuda")
    W = A.clone()
    ms = timeit(lambda: (W.copy_(A), dev.getrf(n, n, W, n, ipiv, info)))
    print(f"zgetrf n={n}: {ms:.2f} ms {(8/3)*n**3/ms*1e-9:.2f} TFLOP/s", flush=True)
    S = (A + A.T.conj()) / 2 + n * torch.eye(n, dtype=torch.complex128, device="cuda")
    ms = timeit(lambda: (W.copy_(S), dev.potrf(dev.LOWER, n, W, n, info)))
    print(f"zpotrf n={n}: {ms:.2f} ms {(4/3)*n**3/ms*1e-9:.2f} TFLOP/s info={info.item()}", flush=True)
    Af = torch.rand((n, n), dtype=torch.float32, device="cuda"); Bf = torch.rand((n, n), dtype=torch.float32, device="cuda"); Cf = torch.zeros((n, n), dtype=torch.float32, device="cuda")
    ms = timeit(lambda: dev.gemm(dev.NO_TRANS, dev.NO_TRANS, n, n, n, 1.0, Af, n, Bf, n, 0.0, Cf, n))
    print(f"sgemm n={n}: {ms:.2f} ms {2*n**3/ms*1e-9:.2f} TFLOP/s path={dev.last_gemm_path()}", flush=True)
