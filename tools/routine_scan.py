"""Timing scan of the LAPACK-level routines through the drop-in symbols (host buffers) for
s / d / z at one moderate size: looks for routines that are far off their neighbours."""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from linalg_b200 import la  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
rng = np.random.default_rng(1)


def tm(fn, reps=2):
    fn()
    best = 1e9
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); best = min(best, time.perf_counter() - t0)
    return best * 1e3


for dt, L in ((np.float32, "s"), (np.float64, "d"), (np.complex128, "z")):
    A = rng.random((n, n)) - 0.5
    if L == "z":
        A = A + 1j * (rng.random((n, n)) - 0.5)
    A = la.GeneralMatrix(np.asfortranarray(A.astype(dt)))
    S = la.GeneralMatrix(np.asfortranarray((A.raw @ A.raw.conj().T / n + np.eye(n)).astype(dt)))
    S.assume_(la.MatrixFlags.PositiveDefinite | la.MatrixFlags.Hermitian | la.MatrixFlags.Symmetric)
    B = la.GeneralMatrix(np.asfortranarray((rng.random((n, 64)) - 0.5).astype(dt)))
    T = la.GeneralMatrix(np.asfortranarray(np.triu(A.raw) + n * np.eye(n, dtype=dt)), flags=la.MatrixFlags.UpperTriangular)
    rows = {
        "gemm": lambda: A * A,
        "lu_factor": lambda: A.clone().lu_factor_() if hasattr(A, "lu_factor_") else A.lu(),
        "solve(64 rhs)": lambda: A.solve(B),
        "inv (getri)": lambda: A.inv(),
        "cholesky": lambda: S.cholesky(),
        "inv PD (potri)": lambda: S.inv(),
        "inv tri (trtri)": lambda: T.inv(),
        "solve tri": lambda: T.solve(B),
        "qr_raw": lambda: A.qr_raw(),
        "qr (orgqr)": lambda: A.qr(),
        "qr pivoted": (lambda: A.qr_raw(pivoting=True)) if n <= 2048 else None,
        "solvels": lambda: A.solvels(B),
        "norm F": lambda: A.norm(),
    }
    for name, fn in rows.items():
        if fn is None:
            continue
        try:
            print(f"{L} n={n} {name:18s} {tm(fn):9.1f} ms", flush=True)
        except Exception as e:  # noqa: BLE001
            print(f"{L} n={n} {name:18s} failed: {type(e).__name__}: {e}", flush=True)
