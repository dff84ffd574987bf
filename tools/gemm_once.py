"""One device-resident DGEMM (Matrix#* pattern) for ncu captures (development tool)."""
import sys
import torch
sys.path.insert(0, ".")
from linalg_b200 import dev
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
torch.manual_seed(0)
A = torch.rand((n, n), dtype=torch.float64, device="cuda")
B = torch.rand((n, n), dtype=torch.float64, device="cuda")
C = torch.zeros((n, n), dtype=torch.float64, device="cuda")
for _ in range(4):
    dev.gemm(dev.NO_TRANS, dev.NO_TRANS, n, n, n, 1.0, A, n, B, n, 1.0, C, n)
torch.cuda.synchronize()
print("ok", dev.last_gemm_path())
