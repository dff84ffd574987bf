"""One device-resident dgels 2^20 x 512 (for ncu launch lists)."""
import sys
import torch
sys.path.insert(0, ".")
from linalg_b200 import dev
m, n = 1 << 20, 512
torch.manual_seed(0)
A0 = torch.rand((n, m), dtype=torch.float64, device="cuda")
b0 = torch.rand((1, m), dtype=torch.float64, device="cuda")
info = torch.zeros(1, dtype=torch.int32, device="cuda")
for rep in range(2):
    A = A0.clone(); b = b0.clone()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    dev.gels(dev.NO_TRANS, m, n, 1, A, m, b, m, info)
    e1.record(); torch.cuda.synchronize()
    print("dgels", e0.elapsed_time(e1), "ms", flush=True)
