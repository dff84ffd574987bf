"""Full residual of a device-resident factorisation at large N (development tool; torch's own
FP64 matmul is only the checker here)."""
import sys
import torch
sys.path.insert(0, ".")
from linalg_b200 import dev  # noqa: E402

which, n = sys.argv[1], int(sys.argv[2])
torch.manual_seed(0)
info = torch.zeros(1, dtype=torch.int32, device="cuda")
if which == "getrf":
    A0 = torch.rand((n, n), dtype=torch.float64, device="cuda")       # (cols, rows): A0[c, r]
    W = A0.clone()
    ipiv = torch.zeros(n, dtype=torch.int32, device="cuda")
    dev.getrf(n, n, W, n, ipiv, info)
    torch.cuda.synchronize()
    M = W.T                                   # math orientation (rows, cols) view
    L = torch.tril(M, -1) + torch.eye(n, dtype=torch.float64, device="cuda")
    U = torch.triu(M)
    LU = L @ U
    # apply the interchanges to A
    PA = A0.T.clone()
    piv = (ipiv.cpu().numpy() - 1)
    perm = list(range(n))
    for i, p in enumerate(piv):
        perm[i], perm[p] = perm[p], perm[i]
    PA = PA[torch.tensor(perm, device="cuda")]
    R = (LU - PA).abs()
else:
    G = torch.rand((n, n), dtype=torch.float64, device="cuda") - 0.5
    S = (G + G.T) / 2 + n * torch.eye(n, dtype=torch.float64, device="cuda")
    W = S.clone()
    dev.potrf(dev.LOWER, n, W, n, info)
    torch.cuda.synchronize()
    Lm = torch.tril(W.T)
    R = (Lm @ Lm.T - S).abs()
    PA = S
print(which, n, "info", info.item(), "max abs residual", R.max().item(), "rel", (R.max() / PA.abs().max()).item())
# locate the worst 512-blocks
nb = 512
blk = R.reshape(n // nb, nb, n // nb, nb).amax(dim=(1, 3))
worst = torch.topk(blk.flatten(), 8)
for v, idx in zip(worst.values.tolist(), worst.indices.tolist()):
    print("  block row", idx // (n // nb), "col", idx % (n // nb), "max", v)
