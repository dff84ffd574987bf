"""Rank / row-space analysis of the intermittent DGEMM error (development tool)."""
import sys
import torch
sys.path.insert(0, ".")
from linalg_b200 import dev  # noqa: E402

torch.manual_seed(5)
n, k = 15360, 512
A = torch.rand((k, n), dtype=torch.float64, device="cuda") - 0.5
C0 = torch.rand((n, n), dtype=torch.float64, device="cuda") - 0.5
Am = A.T
ref = (C0.T - Am @ Am.T).T
low = torch.tril(torch.ones((n, n), dtype=torch.bool, device="cuda")).T
cells = []
for rep in range(120):
    C = C0.clone()
    dev.herk(dev.LOWER, dev.NO_TRANS, n, k, -1.0, A, n, 1.0, C, n)
    E = ((C - ref) * low).T
    bad = (E.abs() > 1e-10).nonzero()
    if len(bad) == 0:
        continue
    cells = sorted(set((int(r) // 8, int(c) // 32) for r, c in bad.tolist()))
    print(f"rep {rep}: {len(bad)} bad elements, {len(cells)} 8x32 cells")
    for cr, cc in cells[:5]:
        rows, cols = slice(cr * 8, cr * 8 + 8), slice(cc * 32, cc * 32 + 32)
        if cr * 8 < cc * 32 + 32:      # touches the diagonal: masked, skip
            continue
        Ec = E[rows, cols]
        sv = torch.linalg.svdvals(Ec)
        rank = int((sv > 1e-9).sum())
        # column-space test: E = A_rows(8 x K) * X  -> is each COLUMN of E in span of A[rows, K]?  (B fragment wrong)
        # row-space test:    E = Y * Bm(K x 32)     -> is each ROW of E in span of Am[cols, K]^T?   (A fragment wrong)
        bestA = bestB = None
        for kb in range(k // 16):
            for ks in range(4):
                ksel = torch.tensor([16 * kb + (t & 1) + 4 * (t >> 1) + 2 * (ks & 1) + 8 * (ks >> 1) for t in range(4)], device="cuda")
                Bk = Am[cols][:, ksel].T                    # 4 x 32
                Y = torch.linalg.lstsq(Bk.T, Ec.T).solution  # (4 x 8): Ec^T ~ Bk^T Y
                rA = (Ec.T - Bk.T @ Y).abs().max().item()
                if bestA is None or rA < bestA[0]:
                    bestA = (rA, kb, ks)
                Ak = Am[rows][:, ksel]                       # 8 x 4
                X = torch.linalg.lstsq(Ak, Ec).solution
                rB = (Ec - Ak @ X).abs().max().item()
                if bestB is None or rB < bestB[0]:
                    bestB = (rB, kb, ks)
        print(f"  cell rows {cr*8} (in-tile {cr*8%128}) cols {cc*32} (in-tile {cc*32%128}): rank {rank} sv {sv[:6].tolist()}")
        print(f"     A-fragment-wrong fit: resid {bestA[0]:.2e} (kb {bestA[1]}, ks {bestA[2]});  B-fragment-wrong fit: resid {bestB[0]:.2e} (kb {bestB[1]}, ks {bestB[2]})")
    break

# ---- recover the wrong A-fragment values of the first full cell and look for them in A
for cr, cc in cells[:5]:
    rows, cols = slice(cr * 8, cr * 8 + 8), slice(cc * 32, cc * 32 + 32)
    if cr * 8 < cc * 32 + 32:
        continue
    Ec = E[rows, cols]
    best = None
    for kb in range(k // 16):
        for ks in range(4):
            kl = [16 * kb + (t & 1) + 4 * (t >> 1) + 2 * (ks & 1) + 8 * (ks >> 1) for t in range(4)]
            ksel = torch.tensor(kl, device="cuda")
            Bk = Am[cols][:, ksel].T
            Y = torch.linalg.lstsq(Bk.T, Ec.T).solution
            r = (Ec.T - Bk.T @ Y).abs().max().item()
            if best is None or r < best[0]:
                best = (r, kl, Y)
    _, kl, Y = best
    # C = C0 - A A^T: E = -(a_wrong - a) b  ->  a_wrong = a - Y^T
    for t, kk in enumerate(kl):
        da = -Y[t]                                   # (8,) = a_wrong - a_true for k = kk
        if da.abs().max().item() < 1e-9:
            print(f"  k={kk}: fragment correct")
            continue
        a_true = Am[rows, kk]
        a_wrong = a_true + da
        for g in range(8):
            if abs(da[g].item()) < 1e-9:
                continue
            val = a_wrong[g].item()
            row = cr * 8 + g
            out = []
            for name, v in (("+da", da[g].item()), ("-da", -da[g].item()), ("wrong", val)):
                hit = ((Am - v).abs() < 1e-11).nonzero()[:3].tolist()
                if hit:
                    out.append(f"{name}={v:+.6f} = A{hit}")
            print(f"  k={kk} row {row} (gid {g}): true {a_true[g].item():+.6f} da {da[g].item():+.6f} {'; '.join(out) if out else 'no exact match for da, -da, wrong'}")
    break
