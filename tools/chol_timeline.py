"""Per-step timeline of the look-ahead Cholesky (needs a -DLA_CHOL_TIMELINE build; lower, FP64)."""
import os
import sys
import torch
sys.path.insert(0, ".")
import linalg_b200
from linalg_b200 import dev
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
G = torch.rand((n, n), dtype=torch.float64, device="cuda") - 0.5
S = (G + G.T) / 2 + n * torch.eye(n, dtype=torch.float64, device="cuda")
info = torch.zeros(1, dtype=torch.int32, device="cuda")
lib = linalg_b200.lib()
lib.la_b200_debug_chol_timeline.restype = None
for rep in range(2):
    W = S.clone()
    torch.cuda.synchronize()
    dev.potrf(dev.LOWER, n, W, n, info)
    torch.cuda.synchronize()
    if rep == 0:   # drop the first (cold) repetition's events
        devnull = os.open(os.devnull, os.O_WRONLY)
        so = os.dup(1); os.dup2(devnull, 1)
        lib.la_b200_debug_chol_timeline()
        os.dup2(so, 1)
lib.la_b200_debug_chol_timeline()
