import sys
import numpy as np
sys.path.insert(0, ".")
import linalg_b200
from linalg_b200 import ffi
abi = linalg_b200.abi()
rng = np.random.default_rng(3)
m, n, k = 2200, 2500, 1600
for tb in (ffi.NO_TRANS, ffi.TRANS):
    A = np.asfortranarray(rng.random((m, k)) - 0.5)
    B = np.asfortranarray(rng.random((k, n) if tb == ffi.NO_TRANS else (n, k)) - 0.5)
    C0 = np.asfortranarray(rng.random((m, n)) - 0.5)
    ref = A @ (B if tb == ffi.NO_TRANS else B.T) + C0
    for rep in range(4):
        C = C0.copy(order="F")
        abi.gemm("d")(ffi.COL_MAJOR, ffi.NO_TRANS, tb, m, n, k, 1.0, ffi.ptr(A), m, ffi.ptr(B), B.shape[0], 1.0, ffi.ptr(C), m)
        bad = np.argwhere(np.abs(C - ref) > 1e-9)
        print(tb, rep, len(bad), bad[:3].tolist(), bad[-3:].tolist() if len(bad) else "")
