"""The drop-in cblas_dgemm on several GPUs of one process (LA_B200_GPUS): result check on
sampled columns against numpy and wall-clock time, host buffers in and out.
  LA_B200_GPUS=2 python tools/multi_gpu_gemm_check.py 16384"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import linalg_b200  # noqa: E402
from linalg_b200 import la  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
rng = np.random.default_rng(5)
A = la.GeneralMatrix(np.asfortranarray(rng.random((n, n))))
B = la.GeneralMatrix(np.asfortranarray(rng.random((n, n))))
with la.use_backend(linalg_b200.abi()):
    for rep in range(3):
        t0 = time.perf_counter()
        C = A * B
        dt = time.perf_counter() - t0
        print(f"LA_B200_GPUS={os.environ.get('LA_B200_GPUS', '1')} cblas_dgemm N={n} rep {rep}: "
              f"{dt * 1e3:.1f} ms = {2.0 * n ** 3 / dt * 1e-12:.1f} TFLOP/s (host in, host out)", flush=True)
cols = [0, 1, n // 2 - 1, n // 2, n // 2 + 129, n - 1]
ref = A.raw @ B.raw[:, cols]
err = np.abs(C.raw[:, cols] - ref).max() / np.abs(ref).max()
print("max relative error on sampled columns", err)
assert err < 1e-13
