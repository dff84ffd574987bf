"""Matrix exponential: device-resident chain against the pointer path (every
product through cblas_dgemm on host buffers) and scipy's expm on the host cores."""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from linalg_b200 import la  # noqa: E402

for n in [int(a) for a in sys.argv[1:]] or [1024, 4096]:
    rng = np.random.default_rng(n)
    A = la.GeneralMatrix(np.asfortranarray((rng.random((n, n)) - 0.5) * (8.0 / np.sqrt(n))))
    A._expm_resident()                     # warm-up (library load, workspaces)
    t0 = time.perf_counter(); r = A._expm_resident(); t1 = time.perf_counter()
    h = A._expm_host(); t2 = time.perf_counter()
    line = f"expm n={n}: resident {1e3 * (t1 - t0):.1f} ms, pointer path {1e3 * (t2 - t1):.1f} ms, " \
           f"difference {np.linalg.norm(r - h) / np.linalg.norm(h):.1e}"
    try:
        import scipy.linalg as sl
        t3 = time.perf_counter(); s = sl.expm(A.raw); t4 = time.perf_counter()
        line += f", scipy (host) {1e3 * (t4 - t3):.1f} ms, difference {np.linalg.norm(r - s) / np.linalg.norm(s):.1e}"
    except ImportError:
        pass
    print(line, flush=True)
