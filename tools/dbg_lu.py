import sys, torch
sys.path.insert(0, ".")
from linalg_b200 import dev
torch.manual_seed(0)
info = torch.zeros(1, dtype=torch.int32, device="cuda")
for n in [int(a) for a in sys.argv[1:]]:
    A0 = torch.rand((n, n), dtype=torch.float64, device="cuda"); W = A0.clone()
    ipiv = torch.zeros(n, dtype=torch.int32, device="cuda")
    for rep in range(2):
        W.copy_(A0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); dev.getrf(n, n, W, n, ipiv, info); e1.record(); torch.cuda.synchronize()
        print("getrf", n, rep, f"{e0.elapsed_time(e1):.2f} ms", flush=True)
