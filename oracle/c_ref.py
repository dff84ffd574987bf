"""TEST INFRASTRUCTURE.  numpy front end of the plain-C oracle (oracle/la_ref.c)."""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

_DIR = Path(__file__).resolve().parent
_SO = _DIR / "_build" / "libla_ref.so"
_lib = None
_LET = {np.dtype(np.float32): "s", np.dtype(np.float64): "d", np.dtype(np.complex128): "z"}


def build():
    subprocess.run(["make", "-C", str(_DIR)], check=True, capture_output=True)
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not _SO.exists():
            build()
        _lib = C.CDLL(str(_SO))
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _f(name, a):
    return getattr(lib(), f"ref_{_LET[a.dtype]}{name}")


def _F(a, dtype=None):
    return np.array(a, dtype=dtype, order="F", copy=True)


def _sc(a, v):
    return np.array([v], dtype=a.dtype)


def gemm(ta, tb, alpha, A, B, beta, Cm):
    """C = alpha*op(A)*op(B) + beta*C; returns the new C."""
    A, B, Cm = _F(A), _F(B, A.dtype), _F(Cm, A.dtype)
    m, n = Cm.shape
    k = A.shape[1] if ta == "N" else A.shape[0]
    al, be = _sc(A, alpha), _sc(A, beta)
    _f("gemm", A)(ord(ta), ord(tb), m, n, k, _p(al), _p(A), A.shape[0], _p(B), B.shape[0], _p(be),
                  _p(Cm), m)
    return Cm


def trsm(side, uplo, trans, diag, alpha, A, B):
    A, B = _F(A), _F(B, np.asarray(A).dtype)
    m, n = B.shape
    al = _sc(A, alpha)
    _f("trsm", A)(ord(side), ord(uplo), ord(trans), ord(diag), m, n, _p(al), _p(A), A.shape[0],
                  _p(B), m)
    return B


def getrf(A):
    A = _F(A)
    m, n = A.shape
    ipiv = np.zeros(min(m, n), dtype=np.int32)
    info = C.c_int(0)
    _f("getrf", A)(m, n, _p(A), m, _p(ipiv), C.byref(info))
    return A, ipiv, info.value


def getrs(trans, LU, ipiv, B):
    LU, B = _F(LU), _F(B, np.asarray(LU).dtype)
    n, nrhs = B.shape
    ipiv = np.ascontiguousarray(ipiv, dtype=np.int32)
    _f("getrs", LU)(ord(trans), n, nrhs, _p(LU), n, _p(ipiv), _p(B), n)
    return B


def potrf(uplo, A):
    A = _F(A)
    n = A.shape[0]
    info = C.c_int(0)
    _f("potrf", A)(ord(uplo), n, _p(A), n, C.byref(info))
    return A, info.value


def potrs(uplo, Cf, B):
    Cf, B = _F(Cf), _F(B, np.asarray(Cf).dtype)
    n, nrhs = B.shape
    _f("potrs", Cf)(ord(uplo), n, nrhs, _p(Cf), n, _p(B), n)
    return B


def trtrs(uplo, trans, diag, A, B):
    A, B = _F(A), _F(B, np.asarray(A).dtype)
    n, nrhs = B.shape
    info = C.c_int(0)
    _f("trtrs", A)(ord(uplo), ord(trans), ord(diag), n, nrhs, _p(A), n, _p(B), n, C.byref(info))
    return B, info.value


def geqrf(A):
    A = _F(A)
    m, n = A.shape
    tau = np.zeros(min(m, n), dtype=A.dtype)
    _f("geqrf", A)(m, n, _p(A), m, _p(tau))
    return A, tau


def ormqr(side, trans, QR, tau, Cm):
    QR, Cm = _F(QR), _F(Cm, np.asarray(QR).dtype)
    m, n = Cm.shape
    tau = np.ascontiguousarray(tau, dtype=QR.dtype)
    _f("ormqr", QR)(ord(side), ord(trans), m, n, len(tau), _p(QR), QR.shape[0], _p(tau), _p(Cm), m)
    return Cm


def gels(trans, A, B):
    """Returns (A_factored, X) with X the first n (trans N) or m rows of the solution block."""
    A = _F(A)
    m, n = A.shape
    B = np.asarray(B, dtype=A.dtype)
    nrhs = B.shape[1]
    ldb = max(m, n)
    Bw = np.zeros((ldb, nrhs), dtype=A.dtype, order="F")
    Bw[: B.shape[0], :] = B
    info = C.c_int(0)
    _f("gels", A)(ord(trans), m, n, nrhs, _p(A), m, _p(Bw), ldb, C.byref(info))
    return A, Bw, info.value


def getri(LU, ipiv):
    LU = _F(LU)
    n = LU.shape[0]
    ipiv = np.ascontiguousarray(ipiv, dtype=np.int32)
    info = C.c_int(0)
    _f("getri", LU)(n, _p(LU), n, _p(ipiv), C.byref(info))
    return LU, info.value


def potri(uplo, Cf):
    Cf = _F(Cf)
    n = Cf.shape[0]
    info = C.c_int(0)
    _f("potri", Cf)(ord(uplo), n, _p(Cf), n, C.byref(info))
    return Cf, info.value


def trtri(uplo, diag, A):
    A = _F(A)
    n = A.shape[0]
    info = C.c_int(0)
    _f("trtri", A)(ord(uplo), ord(diag), n, _p(A), n, C.byref(info))
    return A, info.value
