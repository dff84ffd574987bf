import sys, torch
sys.path.insert(0, ".")
from linalg_b200 import dev
torch.manual_seed(0)
info = torch.zeros(1, dtype=torch.int32, device="cuda")
for (m, n) in [(8192, 16), (8192, 512), (4096, 4096), (8192, 8192)]:
    A = torch.rand((n, m), dtype=torch.complex128, device="cuda")
    ipiv = torch.zeros(min(m, n), dtype=torch.int32, device="cuda")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); dev.getrf(m, n, A, m, ipiv, info); e1.record(); torch.cuda.synchronize()
    print("zgetrf", m, n, f"{e0.elapsed_time(e1):.1f} ms", "info", info.item(), flush=True)
