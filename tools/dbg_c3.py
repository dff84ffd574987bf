import sys, torch
sys.path.insert(0, ".")
from linalg_b200 import dev
torch.manual_seed(0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
info = torch.zeros(1, dtype=torch.int32, device="cuda")
A0 = torch.rand((n, n), dtype=torch.float64, device="cuda"); W = A0.clone()
ipiv = torch.zeros(n, dtype=torch.int32, device="cuda")
def t(name, fn):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fn(); e1.record(); torch.cuda.synchronize()
    print(name, f"{e0.elapsed_time(e1):.2f} ms", flush=True)
t("getrf", lambda: dev.getrf(n, n, W, n, ipiv, info))
for nrhs in (1, 256):
    X = torch.rand((nrhs, n), dtype=torch.float64, device="cuda")
    t(f"getrs nrhs={nrhs}", lambda: dev.getrs(dev.NO_TRANS, n, nrhs, W, n, ipiv, X, n))
m = 1 << 18
A = torch.rand((64, m), dtype=torch.float64, device="cuda"); b = torch.rand((1, m), dtype=torch.float64, device="cuda")
tau = torch.zeros(64, dtype=torch.float64, device="cuda")
t("geqrf 2^18 x 64", lambda: dev.geqrf(m, 64, A, m, tau))
A = torch.rand((64, m), dtype=torch.float64, device="cuda")
t("gels 2^18 x 64", lambda: dev.gels(dev.NO_TRANS, m, 64, 1, A, m, b, m, info))
