"""Catch an intermittent wrong DGEMM result and characterise it (development tool)."""
import sys
import torch
sys.path.insert(0, ".")
from linalg_b200 import dev  # noqa: E402

torch.manual_seed(2)
ta, tb, m, n, k = "N", sys.argv[1] if len(sys.argv) > 1 else "T", 6000, 6000, 1024
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 40
op = {"N": dev.NO_TRANS, "T": dev.TRANS}
A = torch.rand((k, m), dtype=torch.float64, device="cuda") - 0.5            # m x k col-major
B = (torch.rand((k, n) if tb == "T" else (n, k), dtype=torch.float64, device="cuda") - 0.5)
C0 = torch.rand((n, m), dtype=torch.float64, device="cuda") - 0.5
Am = A.T
Bm = B if tb == "T" else B.T            # math k x n: tb == T -> stored n x k col-major = torch (k, n)
ref = (C0.T - Am @ Bm).T
nbad = 0
for rep in range(reps):
    C = C0.clone()
    dev.gemm(op[ta], op[tb], m, n, k, -1.0, A, m, B, n if tb == "T" else k, 1.0, C, m)
    E = (C - ref).T                      # math (m, n)
    bad = (E.abs() > 1e-10).nonzero()
    if len(bad) == 0:
        continue
    nbad += 1
    r0, r1 = bad[:, 0].min().item(), bad[:, 0].max().item()
    c0, c1 = bad[:, 1].min().item(), bad[:, 1].max().item()
    print(f"rep {rep}: {len(bad)} bad elements, rows {r0}..{r1} cols {c0}..{c1} maxerr {E.abs().max().item():.3e}")
    # distinct 8-row x 8-col cells touched
    cells = set((int(r) // 8, int(c) // 8) for r, c in bad[:2000].tolist())
    print("   cells (row/8, col/8) sample:", sorted(cells)[:12], "count", len(cells))
    r, c = bad[0].tolist()
    e = E[r, c].item()
    contrib = (Am[r].reshape(-1, 16) * Bm[:, c].reshape(-1, 16)).sum(dim=1)   # per 16-k block
    # which single block (or +/- multiple) explains the error?  C = C0 - sum; missing block -> E = +contrib
    d = (contrib - e).abs()
    j = int(d.argmin())
    d2 = (contrib + e).abs()
    j2 = int(d2.argmin())
    print(f"   first bad ({r},{c}) err {e:.6e}; best single missing k-block {j} resid {d[j].item():.2e}; "
          f"best duplicated k-block {j2} resid {d2[j2].item():.2e}")
    if nbad >= 3:
        break
print("failures", nbad, "of", rep + 1)

# ---- signature test on the last failure: stale operand data from the neighbouring ring phase?
if nbad:
    cells = sorted(set((int(r) // 8, int(c) // 32) for r, c in bad.tolist()))
    for (cr, cc) in cells[:4]:
        rows = slice(cr * 8, cr * 8 + 8)
        cols = slice(cc * 32, cc * 32 + 32)
        Ecell = E[rows, cols]
        best = None
        for kb in range(k // 16):
            for ks in range(5):                 # 0..3: one DMMA k-step; 4: the whole k-block
                if ks < 4:
                    ksel = [16 * kb + (t & 1) + 4 * (t >> 1) + 2 * (ks & 1) + 8 * (ks >> 1) for t in range(4)]
                else:
                    ksel = list(range(16 * kb, 16 * kb + 16))
                kt = torch.tensor(ksel, device="cuda")
                for delta in (-64, 64, -16, 16, 0):
                    k2 = kt + delta
                    if k2.min() < 0 or k2.max() >= k:
                        continue
                    if delta == 0:
                        pred = (Am[rows][:, kt]) @ Bm[kt][:, cols]          # k-step skipped entirely
                    else:
                        pred = -(Am[rows][:, k2] - Am[rows][:, kt]) @ Bm[kt][:, cols]
                    d = (Ecell - pred).abs().max().item()
                    if best is None or d < best[0]:
                        best = (d, kb, ks, delta)
        print(f"cell rows {cr*8}.. cols {cc*32}..: max|E| {Ecell.abs().max().item():.3e}; best hypothesis "
              f"resid {best[0]:.2e} (kb {best[1]}, ks {best[2]}, delta {best[3]})")
        # prefix / suffix hypotheses: accumulators of this 8-row group reset or frozen at a k-block
        P = torch.zeros_like(Ecell)
        bestp = None
        full = Am[rows] @ Bm[:, cols]
        for kb in range(k // 16 + 1):
            if kb > 0:
                kt = slice(16 * (kb - 1), 16 * kb)
                P = P + Am[rows][:, kt] @ Bm[kt][:, cols]
            for name, pred in (("+prefix", P), ("-prefix", -P), ("+suffix", full - P), ("-suffix", P - full)):
                d = (Ecell - pred).abs().max().item()
                if bestp is None or d < bestp[0]:
                    bestp = (d, kb, name)
        print(f"      prefix/suffix: best resid {bestp[0]:.2e} at k-block {bestp[1]} ({bestp[2]})")
