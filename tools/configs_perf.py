"""Device-resident timings of the BASELINE.json configurations (c1-c5) on one B200.
Prints one JSON object per line; used to fill profiles/ and BASELINE/DESIGN tables."""
import json
import sys

import torch

sys.path.insert(0, ".")
from linalg_b200 import dev  # noqa: E402


def timeit(fn, reps=3, setup=None):
    best = 1e30
    for i in range(reps + 1):
        if setup:
            setup()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        if i:
            best = min(best, e0.elapsed_time(e1))
    return best


def out(**kw):
    print(json.dumps(kw), flush=True)


def main():
    which = sys.argv[1:] or ["c1", "c2", "c3", "c4", "c5"]
    g = torch.Generator(device="cuda"); g.manual_seed(20261106)
    rnd = lambda shape, dt: torch.rand(shape, dtype=dt, device="cuda", generator=g)
    if "c1" in which or "c2" in which:
        sizes = [1024] if "c2" not in which else [512, 1024, 2048, 4096, 8192, 16384]
        for dt, name, fl in ((torch.float64, "dgemm", 2), (torch.float32, "sgemm", 2), (torch.complex128, "zgemm", 8)):
            for n in sizes:
                if dt == torch.complex128 and n > 8192:
                    continue
                A = torch.rand((n, n), dtype=dt, device="cuda"); B = torch.rand((n, n), dtype=dt, device="cuda")
                C = torch.zeros((n, n), dtype=dt, device="cuda")
                ms = timeit(lambda: dev.gemm(dev.NO_TRANS, dev.NO_TRANS, n, n, n, 1.0, A, n, B, n, 1.0, C, n), setup=C.zero_)
                out(config="c2" if n != 1024 else "c1/c2", routine=name, n=n, ms=ms, tflops=fl * n ** 3 / ms * 1e-9, path=dev.last_gemm_path())
                del A, B, C
    info = torch.zeros(1, dtype=torch.int32, device="cuda")
    if "c3" in which:
        n = 16384
        A0 = rnd((n, n), torch.float64); W = torch.empty_like(A0)
        ipiv = torch.zeros(n, dtype=torch.int32, device="cuda")
        for nrhs in (1, 256):
            B0 = rnd((nrhs, n), torch.float64); X = torch.empty_like(B0)
            def solve():
                dev.getrf(n, n, W, n, ipiv, info)
                dev.getrs(dev.NO_TRANS, n, nrhs, W, n, ipiv, X, n)
            ms = timeit(solve, reps=2, setup=lambda: (W.copy_(A0), X.copy_(B0)))
            # residual ||A x - b|| / (||A|| ||x||), torch views are transposes of col-major data
            r = (X @ A0 - B0)
            res = (r.norm() / (A0.norm() * X.norm())).item()
            out(config="c3", routine="dgesv (getrf+getrs)", n=n, nrhs=nrhs, ms=ms,
                tflops=((2 / 3) * n ** 3 + 2 * n * n * nrhs) / ms * 1e-9, residual=res, info=int(info.item()))
        del A0, W
    if "c4" in which:
        n = 16384
        G = rnd((n, n), torch.float64) - 0.5
        S = (G + G.T) / 2 + n * torch.eye(n, dtype=torch.float64, device="cuda"); del G
        W = torch.empty_like(S)
        for nrhs in (1, 256):
            B0 = rnd((nrhs, n), torch.float64); X = torch.empty_like(B0)
            def solve():
                dev.potrf(dev.LOWER, n, W, n, info)
                dev.potrs(dev.LOWER, n, nrhs, W, n, X, n)
            ms = timeit(solve, reps=2, setup=lambda: (W.copy_(S), X.copy_(B0)))
            res = ((X @ S - B0).norm() / (S.norm() * X.norm())).item()
            out(config="c4 (1 GPU, N=16384)", routine="dposv (potrf+potrs)", n=n, nrhs=nrhs, ms=ms,
                tflops=((1 / 3) * n ** 3 + 2 * n * n * nrhs) / ms * 1e-9, residual=res, info=int(info.item()))
        del S, W
    if "c5" in which:
        for (m, n, dt, name, fl) in ((1 << 20, 512, torch.float64, "dgels", 1), (1 << 18, 512, torch.complex128, "zgels", 4)):
            A0 = torch.rand((n, m), dtype=dt, device="cuda"); W = torch.empty_like(A0)
            b0 = torch.rand((1, m), dtype=dt, device="cuda"); X = torch.empty_like(b0)
            ms = timeit(lambda: dev.gels(dev.NO_TRANS, m, n, 1, W, m, X, m, info), reps=1,
                        setup=lambda: (W.copy_(A0), X.copy_(b0)))
            x = X[0, :n]
            r = A0.T @ x - b0[0]
            nres = ((A0.conj() @ r).norm() / (A0.norm() * r.norm())).item()   # ||A^H r|| / (||A|| ||r||)
            out(config="c5", routine=name, m=m, n=n, nrhs=1, ms=ms,
                tflops=fl * (2 * m * n * n - (2 / 3) * n ** 3) / ms * 1e-9, normal_eq_residual=nres,
                info=int(info.item()))
            del A0, W


if __name__ == "__main__":
    main()
