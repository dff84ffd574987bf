"""linalg_b200 — B200-native backend for konovod/linalg's dense hot path.

The product is `linalg_b200/lib/libla_b200.so`, a C-ABI shared library (CUDA, sm_100a)
exporting the CBLAS / LAPACK symbols the reference binds (include/la_b200.h).  This
package holds

  * `lib()` / `abi()`     – load that library (fails loudly when it is not built; there is
                             no CPU fallback anywhere in the package);
  * `linalg_b200.la`      – host-side mirror of the reference's `LA::Matrix` API for the hot
                             path (same method names, flag-driven routine selection, lwork
                             protocol, info -> LinAlgError);
  * `linalg_b200.dev`     – thin wrappers of the device-resident API for torch tensors;
  * `linalg_b200.dist`    – multi-GPU drivers (one process per GPU, torch.distributed).
"""
from __future__ import annotations

import ctypes
import os
from pathlib import Path

_ROOT = Path(__file__).resolve().parent
LIB_PATH = _ROOT / "lib" / "libla_b200.so"

_cdll = None
_abi = None


class BackendUnavailable(RuntimeError):
    pass


def lib() -> ctypes.CDLL:
    """Load libla_b200.so.  Raises BackendUnavailable if it has not been built."""
    global _cdll
    if _cdll is None:
        path = Path(os.environ.get("LA_B200_LIB", LIB_PATH))
        if not path.exists():
            raise BackendUnavailable(
                f"{path} not found: build it with `python -m linalg_b200.build` "
                "(there is no CPU fallback)")
        _cdll = ctypes.CDLL(str(path), mode=ctypes.RTLD_LOCAL)
    return _cdll


def abi():
    """The drop-in ABI of libla_b200.so as a `BlasLapackABI`."""
    global _abi
    if _abi is None:
        from .ffi import BlasLapackABI
        _abi = BlasLapackABI(lib(), prefix="", name="la_b200")
    return _abi


def version() -> str:
    f = lib().la_b200_version
    f.restype = ctypes.c_char_p
    return f().decode()
