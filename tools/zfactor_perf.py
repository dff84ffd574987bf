"""zgetrf / zpotrf device timings (development tool)."""
import sys
import torch
sys.path.insert(0, ".")
from linalg_b200 import dev
from tools.gemm_perf import timeit
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
torch.manual_seed(0)
A0 = torch.rand((n, n), dtype=torch.float64, device="cuda") + 1j * torch.rand((n, n), dtype=torch.float64, device="cuda")
W = torch.empty_like(A0)
ipiv = torch.zeros(n, dtype=torch.int32, device="cuda")
info = torch.zeros(1, dtype=torch.int32, device="cuda")
tcopy = timeit(lambda: W.copy_(A0))
def lu():
    W.copy_(A0); dev.getrf(n, n, W, n, ipiv, info)
ms = timeit(lu) - tcopy
print(f"zgetrf n={n}: {ms:.2f} ms {(8/3)*n**3/ms*1e-9:.2f} TFLOP/s", flush=True)
S = A0 @ A0.conj().T / n + torch.eye(n, dtype=torch.complex128, device="cuda")
def ch():
    W.copy_(S); dev.potrf(dev.LOWER, n, W, n, info)
ms = timeit(ch) - tcopy
print(f"zpotrf n={n}: {ms:.2f} ms {(4/3)*n**3/ms*1e-9:.2f} TFLOP/s info={info.item()}", flush=True)
