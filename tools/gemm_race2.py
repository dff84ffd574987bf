"""Where do the intermittent wrong cells of the DGEMM fall? (development tool)"""
import sys
import torch
sys.path.insert(0, ".")
from linalg_b200 import dev  # noqa: E402

torch.manual_seed(3)
m, n, k = 8000, 7000, 512
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 60
A = torch.rand((k, m), dtype=torch.float64, device="cuda") - 0.5
B = torch.rand((n, k), dtype=torch.float64, device="cuda") - 0.5     # k x n col-major
C0 = torch.rand((n, m), dtype=torch.float64, device="cuda") - 0.5
ref = (C0.T - A.T @ B.T).T
TM, TN = (m + 127) // 128, (n + 127) // 128
nt = TM * TN
fails = 0
for rep in range(reps):
    C = C0.clone()
    dev.gemm(dev.NO_TRANS, dev.NO_TRANS, m, n, k, -1.0, A, m, B, k, 1.0, C, m)
    E = (C - ref).T
    bad = (E.abs() > 1e-10).nonzero()
    if len(bad) == 0:
        continue
    fails += 1
    cells = sorted(set((int(r) // 8, int(c) // 32) for r, c in bad.tolist()))
    for cr, cc in cells:
        row, col = cr * 8, cc * 32
        tm, tn = row // 128, col // 128
        g = tm // 16
        first_m = g * 16
        gsz = min(TM - first_m, 16)
        t = g * 16 * TN + (tm - first_m) + tn * gsz
        cnt = int(((bad[:, 0] // 8 == cr) & (bad[:, 1] // 32 == cc)).sum())
        print(f"rep {rep}: cell row {row} col {col} ({cnt} bad) tile ({tm},{tn}) t={t} cta={t % 148} seq={t // 148} of {(nt + 147) // 148} "
              f"row-in-tile {row % 128} (wm={row % 128 // 64}, i={row % 64 // 8}) col-in-tile {col % 128} (wn={col % 128 // 32})")
print("failures", fails, "of", reps)
