"""Per-step timeline of the look-ahead LU (needs a -DLA_LU_TIMELINE build)."""
import sys
import torch
sys.path.insert(0, ".")
import linalg_b200
from linalg_b200 import dev
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
A0 = torch.rand((n, n), dtype=torch.float64, device="cuda")
ipiv = torch.zeros(n, dtype=torch.int32, device="cuda")
info = torch.zeros(1, dtype=torch.int32, device="cuda")
lib = linalg_b200.lib()
for rep in range(2):
    W = A0.clone()
    torch.cuda.synchronize()
    dev.getrf(n, n, W, n, ipiv, info)
    torch.cuda.synchronize()
    if rep == 0:
        import ctypes
        # drop the first (cold) repetition's events
        lib.la_b200_debug_lu_timeline.restype = None
        import os
        devnull = os.open(os.devnull, os.O_WRONLY)
        so = os.dup(1); os.dup2(devnull, 1)
        lib.la_b200_debug_lu_timeline()
        os.dup2(so, 1)
lib.la_b200_debug_lu_timeline()
