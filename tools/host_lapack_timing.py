"""End-to-end timings of the host-pointer LAPACK symbols with pinned buffers (development tool)."""
import ctypes as C
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import linalg_b200  # noqa: E402

lib = linalg_b200.lib()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
rng = np.random.default_rng(0)


def pinned(shape):
    t = torch.empty(shape, dtype=torch.float64).pin_memory()
    return t


def t_call(fn, reps=3):
    best = 1e30
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        best = min(best, time.perf_counter() - t0)
    return best * 1e3


ni, one = C.c_int(n), C.c_int(1)
info = C.c_int(0)
A0 = torch.rand((n, n), dtype=torch.float64)
S0 = (A0 + A0.T) / 2 + n * torch.eye(n, dtype=torch.float64)
hA, hB = pinned((n, n)), pinned((1, n))
ipiv = np.zeros(n, dtype=np.int32)


def potrf():
    hA.copy_(S0)
    t0 = time.perf_counter()
    lib.dpotrf_(C.c_char_p(b"L"), C.byref(ni), C.c_void_p(hA.data_ptr()), C.byref(ni), C.byref(info))
    return (time.perf_counter() - t0) * 1e3


def getrf():
    hA.copy_(A0)
    t0 = time.perf_counter()
    lib.dgetrf_(C.byref(ni), C.byref(ni), C.c_void_p(hA.data_ptr()), C.byref(ni),
                ipiv.ctypes.data_as(C.c_void_p), C.byref(info))
    return (time.perf_counter() - t0) * 1e3


def gesv():
    hA.copy_(A0)
    hB.fill_(1.0)
    t0 = time.perf_counter()
    lib.dgesv_(C.byref(ni), C.byref(one), C.c_void_p(hA.data_ptr()), C.byref(ni),
               ipiv.ctypes.data_as(C.c_void_p), C.c_void_p(hB.data_ptr()), C.byref(ni), C.byref(info))
    return (time.perf_counter() - t0) * 1e3


for name, fn, fl in (("dpotrf_", potrf, n ** 3 / 3), ("dgetrf_", getrf, 2 * n ** 3 / 3), ("dgesv_ nrhs=1", gesv, 2 * n ** 3 / 3)):
    fn()
    ms = min(fn() for _ in range(3))
    print(f"{name} n={n} host pointers (pinned): {ms:.1f} ms  {fl / ms * 1e-9:.2f} TFLOP/s  info={info.value}", flush=True)
