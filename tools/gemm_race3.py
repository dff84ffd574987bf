"""Is the intermittent DGEMM error a set of DMMA k-steps missing from the accumulators?"""
import sys
import torch
sys.path.insert(0, ".")
from linalg_b200 import dev  # noqa: E402

torch.manual_seed(4)
n, k = 15360, 512
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 30
A = torch.rand((k, n), dtype=torch.float64, device="cuda") - 0.5       # n x k col-major
C0 = torch.rand((n, n), dtype=torch.float64, device="cuda") - 0.5
Am = A.T                                                               # (n, k)
ref = (C0.T - Am @ Am.T).T
low = torch.tril(torch.ones((n, n), dtype=torch.bool, device="cuda")).T
shown = 0
fails = 0
for rep in range(reps):
    C = C0.clone()
    dev.herk(dev.LOWER, dev.NO_TRANS, n, k, -1.0, A, n, 1.0, C, n)
    E = ((C - ref) * low).T                       # math (row, col)
    bad = (E.abs() > 1e-10).nonzero()
    if len(bad) == 0:
        continue
    fails += 1
    blocks = sorted(set((int(r) // 8, int(c) // 8) for r, c in bad.tolist()))
    print(f"rep {rep}: {len(bad)} bad elements in {len(blocks)} 8x8 blocks")
    for br, bc in blocks[:6]:
        rows, cols = slice(br * 8, br * 8 + 8), slice(bc * 8, bc * 8 + 8)
        Eb = E[rows, cols]
        best = None
        for kb in range(k // 16):
            for ks in range(4):
                ksel = torch.tensor([16 * kb + (t & 1) + 4 * (t >> 1) + 2 * (ks & 1) + 8 * (ks >> 1) for t in range(4)], device="cuda")
                prod = Am[rows][:, ksel] @ Am[cols][:, ksel].T
                for sgn in (1.0, -1.0):
                    d = (Eb - sgn * prod).abs().max().item()
                    if best is None or d < best[0]:
                        best = (d, kb, ks, sgn)
        print(f"   block rows {br*8} cols {bc*8} (in-tile row {br*8%128}, col {bc*8%128}): max|E| {Eb.abs().max().item():.3e} "
              f"best single k-step: resid {best[0]:.2e} kb {best[1]} ks {best[2]} sign {best[3]}")
    shown += 1
    if shown >= 2:
        break
print("failures", fails, "of", rep + 1)
