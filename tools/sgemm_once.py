"""One device-resident FP32 product (3xTF32 tcgen05 kernel) for ncu captures."""
import sys
import torch
sys.path.insert(0, ".")
from linalg_b200 import dev
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
A = torch.rand((n, n), dtype=torch.float32, device="cuda")
B = torch.rand((n, n), dtype=torch.float32, device="cuda")
C = torch.zeros((n, n), dtype=torch.float32, device="cuda")
for _ in range(2):
    dev.gemm(dev.NO_TRANS, dev.NO_TRANS, n, n, n, 1.0, A, n, B, n, 0.0, C, n)
torch.cuda.synchronize()
print("ok", dev.last_gemm_path())
