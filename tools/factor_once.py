"""Run one device-resident factorisation (development tool for ncu launch lists)."""
import sys
import torch
sys.path.insert(0, ".")
from linalg_b200 import dev  # noqa: E402

which, n = sys.argv[1], int(sys.argv[2])
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
torch.manual_seed(0)
ipiv = torch.zeros(n, dtype=torch.int32, device="cuda")
info = torch.zeros(1, dtype=torch.int32, device="cuda")
if which == "getrf":
    A0 = torch.rand((n, n), dtype=torch.float64, device="cuda")
else:
    G = torch.rand((n, n), dtype=torch.float64, device="cuda") - 0.5
    A0 = (G + G.T) / 2 + n * torch.eye(n, dtype=torch.float64, device="cuda")
W = torch.empty_like(A0)
for r in range(reps):
    W.copy_(A0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    if which == "getrf":
        dev.getrf(n, n, W, n, ipiv, info)
    else:
        dev.potrf(dev.LOWER, n, W, n, info)
    e1.record()
    torch.cuda.synchronize()
    print(which, n, "rep", r, f"{e0.elapsed_time(e1):.2f} ms info={info.item()}", flush=True)
