"""One device-resident square dgeqrf (for ncu launch lists)."""
import sys, torch
sys.path.insert(0, ".")
from linalg_b200 import dev
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
A0 = torch.rand((n, n), dtype=torch.float64, device="cuda")
tau = torch.zeros(n, dtype=torch.float64, device="cuda")
for r in range(2):
    A = A0.clone(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); dev.geqrf(n, n, A, n, tau); e1.record(); torch.cuda.synchronize()
    print("dgeqrf", n, e0.elapsed_time(e1), "ms", flush=True)
