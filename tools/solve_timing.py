"""Device timing and check of the small-nrhs solves (development tool; torch is the checker)."""
import sys
import torch
sys.path.insert(0, ".")
from linalg_b200 import dev  # noqa: E402
from tools.op_timing import t  # noqa: E402

torch.manual_seed(0)
info = torch.zeros(1, dtype=torch.int32, device="cuda")
for n in (1000, 4096, 16384):
    A0 = torch.rand((n, n), dtype=torch.float64, device="cuda")
    W = A0.clone()
    ipiv = torch.zeros(n, dtype=torch.int32, device="cuda")
    dev.getrf(n, n, W, n, ipiv, info)
    for nrhs in (1, 4, 8, 16):
        X = torch.rand((nrhs, n), dtype=torch.float64, device="cuda")          # n x nrhs col-major
        Bm = (A0.T @ X.T).T.contiguous()
        for tr, name in ((dev.NO_TRANS, "N"), (dev.TRANS, "T")):
            Bt = Bm.clone() if tr == dev.NO_TRANS else (A0 @ X.T).T.contiguous()
            Bw = Bt.clone()
            dev.getrs(tr, n, nrhs, W, n, ipiv, Bw, n)
            err = ((Bw - X).abs().max() / X.abs().max()).item()
            us = t(lambda: dev.getrs(tr, n, nrhs, W, n, ipiv, Bw, n), reps=5)
            print(f"dgetrs {name} n={n} nrhs={nrhs}: {us:9.1f} us  err {err:.2e}", flush=True)
    G = torch.rand((n, n), dtype=torch.float64, device="cuda") - 0.5
    S = (G + G.T) / 2 + n * torch.eye(n, dtype=torch.float64, device="cuda")
    L = S.clone()
    dev.potrf(dev.LOWER, n, L, n, info)
    U = S.clone()
    dev.potrf(dev.UPPER, n, U, n, info)
    for nrhs in (1, 8):
        X = torch.rand((nrhs, n), dtype=torch.float64, device="cuda")
        Bm = (S @ X.T).T.contiguous()
        for uplo, F, name in ((dev.LOWER, L, "L"), (dev.UPPER, U, "U")):
            Bw = Bm.clone()
            dev.potrs(uplo, n, nrhs, F, n, Bw, n)
            err = ((Bw - X).abs().max() / X.abs().max()).item()
            us = t(lambda: dev.potrs(uplo, n, nrhs, F, n, Bw, n), reps=5)
            print(f"dpotrs {name} n={n} nrhs={nrhs}: {us:9.1f} us  err {err:.2e}", flush=True)
