import sys, torch
sys.path.insert(0, ".")
from linalg_b200 import dev
def timeit(fn, reps=3):
    fn(); torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best
torch.manual_seed(0)
POS = True
for (m, n, k) in [(512, 384, 256), (1000, 777, 333), (2048, 2048, 2048), (8192, 8192, 8192), (16384, 16384, 16384)]:
    for (ta, tb) in [(dev.NO_TRANS, dev.NO_TRANS), (dev.TRANS, dev.NO_TRANS), (dev.NO_TRANS, dev.TRANS)]:
        if m > 4096 and ta != dev.NO_TRANS: continue
        A = torch.rand((k, m) if ta == dev.NO_TRANS else (m, k), dtype=torch.float32, device="cuda")
        B = torch.rand((n, k) if tb == dev.NO_TRANS else (k, n), dtype=torch.float32, device="cuda")
        C = torch.zeros((n, m), dtype=torch.float32, device="cuda")
        ms = timeit(lambda: dev.gemm(ta, tb, m, n, k, 1.0, A, A.shape[1], B, B.shape[1], 0.0, C, m))
        err = float("nan")
        if m <= 8192:
            Am = (A.T if ta == dev.NO_TRANS else A).double(); Bm = (B.T if tb == dev.NO_TRANS else B).double()
            ref = (Am @ Bm).T
            bound = (Am.abs() @ Bm.abs()).T
            err = ((C.double() - ref).abs() / bound).max().item()
            ref32 = (Am.float() @ Bm.float()).T.double()
            err32 = ((ref32 - ref).abs() / bound).max().item()
        print(f"sgemm {m}x{n}x{k} ta={ta} tb={tb}: {ms:.3f} ms {2*m*n*k/ms*1e-9:.1f} TFLOP/s path={dev.last_gemm_path()} max|err|/(|A||B|)={err:.2e} (cuBLAS fp32: {err32 if m<=8192 else float('nan'):.2e})", flush=True)
