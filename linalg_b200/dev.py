"""Device-resident API of libla_b200.so for torch CUDA tensors (include/la_b200.h, Part 3).

torch is plumbing here: it owns device memory and the stream; every function below only
forwards raw device pointers to the C ABI.  Matrices are COLUMN-major: a torch tensor `t` of
shape (cols, rows), contiguous, is the column-major matrix rows x cols with ld = rows (use
`colmajor()` / `as_matrix()` to convert).
"""
from __future__ import annotations

import ctypes as C

import torch

from . import lib
from .ffi import CONJ_TRANS, LEFT, LOWER, NO_TRANS, NON_UNIT, RIGHT, TRANS, UNIT, UPPER  # noqa: F401

_LET = {torch.float32: "s", torch.float64: "d", torch.complex128: "z"}
_i64, _p, _int = C.c_int64, C.c_void_p, C.c_int


def _stream():
    return _p(torch.cuda.current_stream().cuda_stream)


_FN_CACHE = {}


def _fn(name, argtypes):
    f = _FN_CACHE.get(name)
    if f is None:
        f = getattr(lib(), name)
        f.argtypes = argtypes
        f.restype = _int
        _FN_CACHE[name] = f
    return f


def _check(rc, name):
    if rc != 0:
        raise ValueError(f"{name}: argument {-rc} is invalid")


def colmajor(rows: int, cols: int, dtype=torch.float64, device="cuda"):
    """Uninitialised column-major rows x cols matrix (a (cols, rows) torch tensor)."""
    return torch.empty((cols, rows), dtype=dtype, device=device)


def _scalar(letter, v):
    if letter == "z":
        import numpy as np
        buf = np.array([complex(v)], dtype=np.complex128)
        return buf.ctypes.data_as(_p), buf
    return (C.c_float(v) if letter == "s" else C.c_double(v)), None


def gemm(ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, Cm, ldc):
    L = _LET[A.dtype]
    al, k1 = _scalar(L, alpha)
    be, k2 = _scalar(L, beta)
    sc = _p if L == "z" else (C.c_float if L == "s" else C.c_double)
    f = _fn(f"la_b200_{L}gemm_dev", [_int, _int, _i64, _i64, _i64, sc, _p, _i64, _p, _i64, sc, _p,
                                     _i64, _p])
    _check(f(ta, tb, m, n, k, al, A.data_ptr(), lda, B.data_ptr(), ldb, be, Cm.data_ptr(), ldc,
             _stream()), "gemm")


def gemm_stair(ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, Cm, ldc, nb, row_first, row_step,
               col_first, col_step):
    """Product restricted to the blocks on/below a block staircase (la_b200_Xgemm_stair_dev)."""
    L = _LET[A.dtype]
    al, k1 = _scalar(L, alpha)
    be, k2 = _scalar(L, beta)
    sc = _p if L == "z" else (C.c_float if L == "s" else C.c_double)
    f = _fn(f"la_b200_{L}gemm_stair_dev", [_int, _int, _i64, _i64, _i64, sc, _p, _i64, _p, _i64, sc,
                                           _p, _i64, _i64, _i64, _i64, _i64, _i64, _p])
    _check(f(ta, tb, m, n, k, al, A.data_ptr(), lda, B.data_ptr(), ldb, be, Cm.data_ptr(), ldc, nb,
             row_first, row_step, col_first, col_step, _stream()), "gemm_stair")


def set_sm_limit(sms: int) -> int:
    """Persistent kernels of this thread's later calls use at most `sms` SMs (0 = all)."""
    f = lib().la_b200_set_sm_limit
    f.argtypes = [_int]
    f.restype = _int
    return int(f(int(sms)))


def laswp(m, n, A, lda, k1, k2, ipiv):
    L = _LET[A.dtype]
    f = _fn(f"la_b200_{L}laswp_dev", [_i64, _i64, _p, _i64, _i64, _i64, _p, _p])
    _check(f(m, n, A.data_ptr(), lda, k1, k2, ipiv.data_ptr(), _stream()), "laswp")


def getrf(m, n, A, lda, ipiv, info):
    L = _LET[A.dtype]
    f = _fn(f"la_b200_{L}getrf_dev", [_i64, _i64, _p, _i64, _p, _p, _p])
    _check(f(m, n, A.data_ptr(), lda, ipiv.data_ptr(), info.data_ptr(), _stream()), "getrf")


def getrs(trans, n, nrhs, A, lda, ipiv, B, ldb):
    L = _LET[A.dtype]
    f = _fn(f"la_b200_{L}getrs_dev", [_int, _i64, _i64, _p, _i64, _p, _p, _i64, _p])
    _check(f(trans, n, nrhs, A.data_ptr(), lda, ipiv.data_ptr(), B.data_ptr(), ldb, _stream()),
           "getrs")


def potrf(uplo, n, A, lda, info):
    L = _LET[A.dtype]
    f = _fn(f"la_b200_{L}potrf_dev", [_int, _i64, _p, _i64, _p, _p])
    _check(f(uplo, n, A.data_ptr(), lda, info.data_ptr(), _stream()), "potrf")


def potrs(uplo, n, nrhs, A, lda, B, ldb):
    L = _LET[A.dtype]
    f = _fn(f"la_b200_{L}potrs_dev", [_int, _i64, _i64, _p, _i64, _p, _i64, _p])
    _check(f(uplo, n, nrhs, A.data_ptr(), lda, B.data_ptr(), ldb, _stream()), "potrs")


def trsm(side, uplo, trans, diag, m, n, alpha, A, lda, B, ldb):
    L = _LET[A.dtype]
    import numpy as np
    dt = {"s": np.float32, "d": np.float64, "z": np.complex128}[L]
    buf = np.array([alpha], dtype=dt)
    f = _fn(f"la_b200_{L}trsm_dev", [_int, _int, _int, _int, _i64, _i64, _p, _p, _i64, _p, _i64, _p])
    _check(f(side, uplo, trans, diag, m, n, buf.ctypes.data_as(_p), A.data_ptr(), lda,
             B.data_ptr(), ldb, _stream()), "trsm")


def herk(uplo, trans, n, k, alpha, A, lda, beta, Cm, ldc):
    L = _LET[A.dtype]
    f = _fn(f"la_b200_{L}herk_dev", [_int, _int, _i64, _i64, C.c_double, _p, _i64, C.c_double, _p,
                                     _i64, _p])
    _check(f(uplo, trans, n, k, alpha, A.data_ptr(), lda, beta, Cm.data_ptr(), ldc, _stream()),
           "herk")


def geqrf(m, n, A, lda, tau):
    L = _LET[A.dtype]
    f = _fn(f"la_b200_{L}geqrf_dev", [_i64, _i64, _p, _i64, _p, _p])
    _check(f(m, n, A.data_ptr(), lda, tau.data_ptr(), _stream()), "geqrf")


def ormqr(side, trans, m, n, k, A, lda, tau, Cm, ldc):
    L = _LET[A.dtype]
    f = _fn(f"la_b200_{L}ormqr_dev", [_int, _int, _i64, _i64, _i64, _p, _i64, _p, _p, _i64, _p])
    _check(f(side, trans, m, n, k, A.data_ptr(), lda, tau.data_ptr(), Cm.data_ptr(), ldc,
             _stream()), "ormqr")


def gels(trans, m, n, nrhs, A, lda, B, ldb, info):
    L = _LET[A.dtype]
    f = _fn(f"la_b200_{L}gels_dev", [_int, _i64, _i64, _i64, _p, _i64, _p, _i64, _p, _p])
    _check(f(trans, m, n, nrhs, A.data_ptr(), lda, B.data_ptr(), ldb, info.data_ptr(), _stream()),
           "gels")


def trtrs(uplo, trans, diag, n, nrhs, A, lda, B, ldb, info):
    L = _LET[A.dtype]
    f = _fn(f"la_b200_{L}trtrs_dev", [_int, _int, _int, _i64, _i64, _p, _i64, _p, _i64, _p, _p])
    _check(f(uplo, trans, diag, n, nrhs, A.data_ptr(), lda, B.data_ptr(), ldb, info.data_ptr(),
             _stream()), "trtrs")


def last_gemm_path() -> int:
    f = lib().la_b200_last_gemm_path
    f.restype = _int
    return int(f())
