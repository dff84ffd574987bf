"""Correctness stress of the FP64 tensor-core GEMM on the shapes the factorisations use
(development tool; torch's FP64 matmul is the checker)."""
import sys
import torch
sys.path.insert(0, ".")
from linalg_b200 import dev  # noqa: E402

torch.manual_seed(1)
SHAPES = [("N", "T", 7040, 896, 128), ("N", "T", 7168, 1024, 1024), ("N", "N", 15872, 512, 512),
          ("N", "N", 8000, 7000, 512), ("N", "T", 6000, 6000, 1024), ("N", "N", 16128, 128, 128),
          ("N", "T", 16256, 64, 64), ("N", "N", 3000, 3000, 64), ("N", "N", 5000, 4000, 32),
          ("T", "N", 4096, 4096, 256), ("N", "T", 3333, 2777, 200)]
op = {"N": dev.NO_TRANS, "T": dev.TRANS}
bad = 0
for rep in range(3):
    for ta, tb, m, n, k in SHAPES:
        ar, ac = (m, k) if ta == "N" else (k, m)
        br, bc = (k, n) if tb == "N" else (n, k)
        A = torch.rand((ac, ar), dtype=torch.float64, device="cuda") - 0.5
        B = torch.rand((bc, br), dtype=torch.float64, device="cuda") - 0.5
        C0 = torch.rand((n, m), dtype=torch.float64, device="cuda") - 0.5
        C = C0.clone()
        dev.gemm(op[ta], op[tb], m, n, k, -1.0, A, ar, B, br, 1.0, C, m)
        Am = A.T if ta == "N" else A          # math (m x k)
        Bm = B.T if tb == "N" else B          # math (k x n)
        ref = (C0.T - Am @ Bm).T
        err = (C - ref).abs().max().item()
        tol = 1e-13 * k
        flag = "" if err < tol else "  <-- WRONG"
        bad += err >= tol
        if rep == 0 or flag:
            print(f"rep {rep} gemm {ta}{tb} {m}x{n}x{k}: max err {err:.3e}{flag}", flush=True)
    # masked (SYRK-style) update, lower
    for n, k in ((7168, 1024), (3000, 128), (15360, 512)):
        A = torch.rand((k, n), dtype=torch.float64, device="cuda") - 0.5
        C0 = torch.rand((n, n), dtype=torch.float64, device="cuda") - 0.5
        C = C0.clone()
        dev.herk(dev.LOWER, dev.NO_TRANS, n, k, -1.0, A, n, 1.0, C, n)
        ref = (C0.T - A.T @ A).T
        low = torch.tril(torch.ones((n, n), dtype=torch.bool, device="cuda")).T   # torch view of lower
        err = ((C - ref).abs() * low).max().item()
        untouched = ((C - C0).abs() * (~low)).max().item()
        flag = "" if (err < 1e-13 * k and untouched == 0.0) else "  <-- WRONG"
        bad += bool(flag)
        if rep == 0 or flag:
            print(f"rep {rep} herk L {n} k={k}: max err {err:.3e} upper touched {untouched:.1e}{flag}", flush=True)
print("bad", bad)
