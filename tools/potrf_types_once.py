"""One device-resident ?potrf / ?getrf of another element type (for ncu launch lists)."""
import sys, torch
sys.path.insert(0, ".")
from linalg_b200 import dev
which, n, dt = sys.argv[1], int(sys.argv[2]), {"s": torch.float32, "d": torch.float64, "z": torch.complex128}[sys.argv[3]]
G = torch.rand((n, n), dtype=dt, device="cuda") - 0.5
A0 = G if which == "getrf" else (G @ G.mH) / n + torch.eye(n, dtype=dt, device="cuda")
ipiv = torch.zeros(n, dtype=torch.int32, device="cuda"); info = torch.zeros(1, dtype=torch.int32, device="cuda")
for r in range(2):
    W = A0.clone(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    if which == "getrf": dev.getrf(n, n, W, n, ipiv, info)
    else: dev.potrf(dev.LOWER, n, W, n, info)
    e1.record(); torch.cuda.synchronize()
    print(which, sys.argv[3], n, e0.elapsed_time(e1), "ms info", info.item(), flush=True)
