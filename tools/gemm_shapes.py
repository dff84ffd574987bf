"""Device timing of the GEMM shapes that occur inside the factorisations (development tool)."""
import sys

import torch

sys.path.insert(0, ".")
from linalg_b200 import dev  # noqa: E402
from tools.gemm_perf import timeit  # noqa: E402

SHAPES = [  # (name, ta, tb, m, n, k)
    ("NN square", "N", "N", 16384, 16384, 16384 // 4),
    ("NT chol trailing full", "N", "T", 15360, 15360, 1024),
    ("NN lu trailing k=512", "N", "N", 15872, 15872, 512),
    ("NN lu trailing k=256", "N", "N", 15872, 15872, 256),
    ("NN lu trailing k=128", "N", "N", 15872, 15872, 128),
    ("NT chol panel k=128", "N", "T", 16256, 896, 128),
    ("NT chol panel k=128 small", "N", "T", 8192, 512, 128),
    ("NN lu in-panel k=128", "N", "N", 16128, 128, 128),
    ("NN lu in-panel k=64", "N", "N", 16128, 64, 64),
    ("NT trsm k=64", "N", "T", 16256, 64, 64),
]


def main():
    torch.manual_seed(0)
    op = {"N": dev.NO_TRANS, "T": dev.TRANS}
    for name, ta, tb, m, n, k in SHAPES:
        ar, ac = (m, k) if ta == "N" else (k, m)
        br, bc = (k, n) if tb == "N" else (n, k)
        A = torch.rand((ac, ar), dtype=torch.float64, device="cuda")
        B = torch.rand((bc, br), dtype=torch.float64, device="cuda")
        Cm = torch.zeros((n, m), dtype=torch.float64, device="cuda")
        ms = timeit(lambda: dev.gemm(op[ta], op[tb], m, n, k, -1.0, A, ar, B, br, 1.0, Cm, m), reps=5, warm=2)
        print(f"{name:28s} {ta}{tb} {m}x{n}x{k}: {ms*1e3:9.1f} us {2*m*n*k/ms*1e-9:6.2f} TFLOP/s path={dev.last_gemm_path()}", flush=True)
    n, k = 15360, 1024
    A = torch.rand((k, n), dtype=torch.float64, device="cuda")
    Cm = torch.zeros((n, n), dtype=torch.float64, device="cuda")
    ms = timeit(lambda: dev.herk(dev.LOWER, dev.NO_TRANS, n, k, -1.0, A, n, 1.0, Cm, n), reps=5, warm=2)
    print(f"herk lower {n} k={k}: {ms*1e3:9.1f} us {n*n*k/ms*1e-9:6.2f} TFLOP/s (useful flops)", flush=True)


if __name__ == "__main__":
    main()
