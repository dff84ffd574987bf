"""Per-phase cycle counts of the LU panel kernel (needs a -DLA_PANEL_TRACE build)."""
import ctypes, sys
import numpy as np
import torch
sys.path.insert(0, ".")
import linalg_b200
from linalg_b200 import dev
lib = linalg_b200.lib()
for m in (512, 4096, 16384):
    A = torch.rand((32, m), dtype=torch.float64, device="cuda")
    ipiv = torch.zeros(32, dtype=torch.int32, device="cuda")
    info = torch.zeros(1, dtype=torch.int32, device="cuda")
    for _ in range(3):
        dev.getrf(m, 32, A, m, ipiv, info)
    out = np.zeros(64 * 8, dtype=np.int64)
    lib.la_b200_debug_panel_trace(out.ctypes.data_as(ctypes.c_void_p))
    t = out.reshape(64, 8)[:32]
    # slots: 0 column start, 1 after barrier A (warp candidates), 2 after barrier B (row staged),
    # 7 candidate completed, 3 published, 4 winner known, 5 after barrier 2, 6 head done
    order = [0, 1, 2, 7, 3, 4, 5, 6] if m > 512 else [0, 1, 2, 5, 6]
    names = ["argmax+barrierA", "pick+stage+barrierB", "fixup", "publish", "rinv+tails+gather",
             "fetch+barrier2", "head"] if m > 512 else \
            ["argmax+barrierA", "pick+stage+barrierB", "warp0+tails+barrier2", "head"]
    d = np.diff(t[:, order], axis=1)[2:30].mean(axis=0).round(0)
    nxt = t[1:, 0] - t[:-1, 6]
    print(f"m={m}: cycles per phase", dict(zip(names, d.tolist())), "loop-back", nxt[2:30].mean().round(0),
          "per column", (t[30, 0] - t[2, 0]) / 28)
