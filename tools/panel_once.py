import sys
import torch
sys.path.insert(0, ".")
from linalg_b200 import dev
m = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
torch.manual_seed(0)
info = torch.zeros(1, dtype=torch.int32, device="cuda")
A = torch.rand((32, m), dtype=torch.float64, device="cuda")
ipiv = torch.zeros(32, dtype=torch.int32, device="cuda")
for _ in range(3):
    dev.getrf(m, 32, A, m, ipiv, info)
torch.cuda.synchronize()
print("ok", info.item())
