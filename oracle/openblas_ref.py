"""TEST INFRASTRUCTURE.  The reference's real arithmetic back end, as a `BlasLapackABI`.

konovod/linalg links `-lcblas -llapack` (src/linalg/libCBLAS.cr:1-7,
src/linalg/libLAPACK_linux.cr:5-9); its CI resolves these to OpenBLAS + LAPACK 3.12.0.
That library is not part of /root/reference and is not pinned by shard.yml.  SciPy ships
the same code (OpenBLAS 0.3.31.dev, LP64, LAPACK 3.12.0) as scipy.libs/libscipy_openblas-*.so
with every symbol prefixed `scipy_`; the ABI is otherwise identical.
"""
from __future__ import annotations

import ctypes
import glob
import os

from linalg_b200.ffi import BlasLapackABI

_abi = None


def find_library() -> str:
    import scipy
    pat = os.path.join(os.path.dirname(scipy.__file__), "..", "scipy.libs",
                       "libscipy_openblas-*.so")
    hits = sorted(glob.glob(pat))
    if not hits:
        raise RuntimeError("scipy's bundled OpenBLAS not found")
    return os.path.realpath(hits[0])


def abi() -> BlasLapackABI:
    global _abi
    if _abi is None:
        cdll = ctypes.CDLL(find_library(), mode=ctypes.RTLD_LOCAL)
        _abi = BlasLapackABI(cdll, prefix="scipy_", name="openblas")
    return _abi


def set_num_threads(n: int):
    f = abi().cdll.scipy_openblas_set_num_threads
    f.argtypes = [ctypes.c_int]
    f(int(n))


def get_num_threads() -> int:
    f = abi().cdll.scipy_openblas_get_num_threads
    f.restype = ctypes.c_int
    return int(f())


def config() -> str:
    f = abi().cdll.scipy_openblas_get_config
    f.restype = ctypes.c_char_p
    return f().decode()
