"""Quick device-resident timing of the FP64 GEMM / LU / Cholesky paths (development tool)."""
import sys
import time

import torch

sys.path.insert(0, ".")
from linalg_b200 import dev  # noqa: E402


def timeit(fn, reps=3, warm=1):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


def main():
    sizes = [int(a) for a in sys.argv[1:]] or [1024, 4096, 8192, 16384]
    torch.manual_seed(0)
    for n in sizes:
        A = torch.rand((n, n), dtype=torch.float64, device="cuda")
        B = torch.rand((n, n), dtype=torch.float64, device="cuda")
        Cm = torch.zeros((n, n), dtype=torch.float64, device="cuda")
        for name, (ta, tb) in {"NN": (dev.NO_TRANS, dev.NO_TRANS), "NT": (dev.NO_TRANS, dev.TRANS),
                               "TN": (dev.TRANS, dev.NO_TRANS), "TT": (dev.TRANS, dev.TRANS)}.items():
            ms = timeit(lambda: dev.gemm(ta, tb, n, n, n, 1.0, A, n, B, n, 0.0, Cm, n))
            # torch tensors are row-major views of our column-major matrices: X_cm = X_t^T
            At = A.T if ta == dev.NO_TRANS else A
            Bt = B.T if tb == dev.NO_TRANS else B
            if n <= 4096:
                ref = (At @ Bt).T.contiguous()
                err = ((Cm - ref).norm() / ref.norm()).item()
            else:
                err = float("nan")
            print(f"dgemm {name} n={n}: {ms:.2f} ms  {2*n**3/ms*1e-9:.2f} TFLOP/s  path={dev.last_gemm_path()} relerr={err:.2e}", flush=True)
        # LU and Cholesky
        ipiv = torch.zeros(n, dtype=torch.int32, device="cuda")
        info = torch.zeros(1, dtype=torch.int32, device="cuda")
        A0 = torch.rand((n, n), dtype=torch.float64, device="cuda")
        W = torch.empty_like(A0)

        def lu():
            W.copy_(A0)
            dev.getrf(n, n, W, n, ipiv, info)
        tcopy = timeit(lambda: W.copy_(A0))
        ms = timeit(lu) - tcopy
        print(f"dgetrf n={n}: {ms:.2f} ms  {(2/3)*n**3/ms*1e-9:.2f} TFLOP/s info={info.item()}", flush=True)
        G = torch.rand((n, n), dtype=torch.float64, device="cuda") - 0.5
        S = (G + G.T) / 2 + n * torch.eye(n, dtype=torch.float64, device="cuda")

        def chol():
            W.copy_(S)
            dev.potrf(dev.LOWER, n, W, n, info)
        ms = timeit(chol) - tcopy
        print(f"dpotrf n={n}: {ms:.2f} ms  {(1/3)*n**3/ms*1e-9:.2f} TFLOP/s info={info.item()}", flush=True)
        del A, B, Cm, A0, W, G, S
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
