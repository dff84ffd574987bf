import sys, torch
sys.path.insert(0, ".")
from linalg_b200 import dev
n = int(sys.argv[1]); dt = {"s": torch.float32, "z": torch.complex128}[sys.argv[2]]
A0 = torch.rand((n, n), dtype=dt, device="cuda")
ipiv = torch.zeros(n, dtype=torch.int32, device="cuda"); info = torch.zeros(1, dtype=torch.int32, device="cuda")
for r in range(2):
    W = A0.clone(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); dev.getrf(n, n, W, n, ipiv, info); e1.record(); torch.cuda.synchronize()
    print("getrf", sys.argv[2], n, e0.elapsed_time(e1), "ms", flush=True)
