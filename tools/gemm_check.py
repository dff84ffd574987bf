import sys, torch
sys.path.insert(0, ".")
from linalg_b200 import dev
torch.manual_seed(1)
def chk(ta, tb, m, n, k):
    A = torch.rand((k, m) if ta == dev.NO_TRANS else (m, k), dtype=torch.float64, device="cuda") - 0.5
    B = torch.rand((n, k) if tb == dev.NO_TRANS else (k, n), dtype=torch.float64, device="cuda") - 0.5
    C = torch.rand((n, m), dtype=torch.float64, device="cuda")
    C0 = C.clone()
    lda = A.shape[1]; ldb = B.shape[1]
    dev.gemm(ta, tb, m, n, k, 1.0, A, lda, B, ldb, 1.0, C, m)
    Am = A.T if ta == dev.NO_TRANS else A
    Bm = B.T if tb == dev.NO_TRANS else B
    ref = (Am @ Bm).T + C0
    err = ((C - ref).norm() / ref.norm()).item()
    bad = ((C - ref).abs() > 1e-9).nonzero()
    print(ta, tb, m, n, k, "relerr", err, "path", dev.last_gemm_path(), "nbad", len(bad), bad[:5].tolist() if len(bad) else "")
for (m, n, k) in [(2200, 384, 1600), (2200, 196, 1600), (2200, 2500, 1600), (256, 196, 64), (128, 4, 64), (128, 20, 16)]:
    chk(dev.NO_TRANS, dev.TRANS, m, n, k)
    chk(dev.NO_TRANS, dev.NO_TRANS, m, n, k)
    chk(dev.TRANS, dev.TRANS, m, n, k)
