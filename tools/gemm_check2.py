import sys, torch
sys.path.insert(0, ".")
from linalg_b200 import dev
torch.manual_seed(1)
def chk(ta, tb, m, n, k):
    A = torch.rand((k, m) if ta == dev.NO_TRANS else (m, k), dtype=torch.float64, device="cuda") - 0.5
    B = torch.rand((n, k) if tb == dev.NO_TRANS else (k, n), dtype=torch.float64, device="cuda") - 0.5
    C = torch.zeros((n, m), dtype=torch.float64, device="cuda")
    lda = A.shape[1]; ldb = B.shape[1]
    Am = A.T if ta == dev.NO_TRANS else A
    Bm = B.T if tb == dev.NO_TRANS else B
    ref = (Am @ Bm).T
    res = []
    for rep in range(3):
        C.zero_()
        dev.gemm(ta, tb, m, n, k, 1.0, A, lda, B, ldb, 0.0, C, m)
        bad = ((C - ref).abs() > 1e-9).nonzero()
        res.append((len(bad), bad[0].tolist() if len(bad) else None, bad[-1].tolist() if len(bad) else None))
    print(m, n, k, res)
for shp in [(2200, 2500, 1600), (2176, 2432, 1600), (2200, 2560, 1600), (2304, 2500, 1600), (2304, 2560, 1600), (2304, 2560, 160), (4096, 4096, 512), (2200, 2500, 64), (2200, 2500, 512)]:
    chk(dev.NO_TRANS, dev.TRANS, *shp)
