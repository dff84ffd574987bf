"""Per-phase cycle counts of the shared-memory QR panel kernel (needs a -DLA_PANEL_TRACE build)."""
import ctypes, sys
import numpy as np
import torch
sys.path.insert(0, ".")
import linalg_b200
from linalg_b200 import dev
lib = linalg_b200.lib()
for m in (512, 4096, 16384):
    A = torch.rand((32, m), dtype=torch.float64, device="cuda")
    tau = torch.zeros(32, dtype=torch.float64, device="cuda")
    for _ in range(3):
        W = A.clone()
        dev.geqrf(m, 32, W, m, tau)
    out = np.zeros(64 * 8, dtype=np.int64)
    lib.la_b200_debug_qr_trace(out.ctypes.data_as(ctypes.c_void_p))
    t = out.reshape(64, 8)[:32, :6]
    d = np.diff(t, axis=1)[2:30].mean(axis=0).round(0)
    names = ["sums+publish", "reduce all CTAs", "barrier", "householder+barrier", "apply+barrier"]
    print(f"m={m}: cycles per phase", dict(zip(names, d.tolist())), "per column", (t[30, 0] - t[2, 0]) / 28)
