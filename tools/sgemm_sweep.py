"""FP32 GEMM sweep (3xTF32 tcgen05): device-resident timing and accuracy against FP64."""
import sys
import torch
sys.path.insert(0, ".")
from linalg_b200 import dev

for n in [int(a) for a in sys.argv[1:]] or [512, 1024, 2048, 4096, 8192, 16384]:
    g = torch.Generator(device="cuda").manual_seed(n)
    A = torch.rand((n, n), dtype=torch.float32, device="cuda", generator=g) - 0.5
    B = torch.rand((n, n), dtype=torch.float32, device="cuda", generator=g) - 0.5
    C = torch.zeros((n, n), dtype=torch.float32, device="cuda")
    best = 1e30
    for i in range(4):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        dev.gemm(dev.NO_TRANS, dev.NO_TRANS, n, n, n, 1.0, A, n, B, n, 0.0, C, n)
        e1.record()
        torch.cuda.synchronize()
        if i:
            best = min(best, e0.elapsed_time(e1))
    cols = [0, n // 3, n - 1]
    ref = torch.stack([A.T.double() @ B[c].double() for c in cols])
    err = float(((C[cols].double() - ref).norm() / ref.norm()).item())
    C2 = torch.zeros_like(C)
    dev.gemm(dev.NO_TRANS, dev.NO_TRANS, n, n, n, 1.0, A, n, B, n, 0.0, C2, n)
    print(f"sgemm n={n}: {best:.3f} ms {2 * n ** 3 / best * 1e-9:.1f} TFLOP/s path={dev.last_gemm_path()} "
          f"relerr_vs_f64={err:.2e} repeatable={bool(torch.equal(C, C2))}", flush=True)
