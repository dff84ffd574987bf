"""Per-phase cycle counts of the Cholesky leaf kernel (needs a -DLA_PANEL_TRACE build)."""
import ctypes, sys
import numpy as np
import torch
sys.path.insert(0, ".")
import linalg_b200
from linalg_b200 import dev
lib = linalg_b200.lib()
n = 128
G = torch.rand((n, n), dtype=torch.float64, device="cuda") - 0.5
S = G @ G.T + n * torch.eye(n, dtype=torch.float64, device="cuda")
info = torch.zeros(1, dtype=torch.int32, device="cuda")
for _ in range(3):
    W = S.clone()
    dev.potrf(dev.LOWER, n, W, n, info)
out = np.zeros(32, dtype=np.int64)
lib.la_b200_debug_potf_trace(out.ctypes.data_as(ctypes.c_void_p))
t = out
print("load", t[1] - t[0])
for b in range(4):
    print("block", b, "diag", t[2 + 3 * b] - (t[1] if b == 0 else t[4 + 3 * (b - 1)]),
          "solve", t[3 + 3 * b] - t[2 + 3 * b], "update", t[4 + 3 * b] - t[3 + 3 * b])
print("store", t[21] - t[20], "total", t[21] - t[0])
