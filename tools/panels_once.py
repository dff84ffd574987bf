"""One launch each of the latency-bound panel kernels, for ncu captures (development tool)."""
import sys
import torch
sys.path.insert(0, ".")
from linalg_b200 import dev
torch.manual_seed(0)
info = torch.zeros(1, dtype=torch.int32, device="cuda")
m = 16384
A = torch.rand((32, m), dtype=torch.float64, device="cuda")
ipiv = torch.zeros(32, dtype=torch.int32, device="cuda")
for _ in range(3):
    dev.getrf(m, 32, A, m, ipiv, info)                       # lu_panel_kernel<double, 1>
G = torch.rand((128, 128), dtype=torch.float64, device="cuda") - 0.5
S = G @ G.T + 128 * torch.eye(128, dtype=torch.float64, device="cuda")
for _ in range(3):
    W = S.clone()
    dev.potrf(dev.LOWER, 128, W, 128, info)                  # potf2_kernel<double>
mq = 1 << 20
Q = torch.rand((32, mq), dtype=torch.float64, device="cuda")
tau = torch.zeros(32, dtype=torch.float64, device="cuda")
for _ in range(2):
    dev.geqrf(mq, 32, Q, mq, tau)                            # qr_panel_stream_kernel<double>
n = 16384
T = torch.rand((n, n), dtype=torch.float64, device="cuda") + n * torch.eye(n, dtype=torch.float64, device="cuda")
b = torch.rand((1, n), dtype=torch.float64, device="cuda")
for _ in range(2):
    dev.trtrs(dev.LOWER, dev.NO_TRANS, dev.NON_UNIT, n, 1, T, n, b, n, info)   # trsv_kernel<double>
torch.cuda.synchronize()
print("ok")
