"""TEST INFRASTRUCTURE.  Residual and pivot checks with BASELINE.json's tolerances.

north_star: relative residual ||Ax-b||/(||A|| ||x||) and ||A-PLU||/||A|| <= 1e-12 for FP64
(and complex128), 1e-5 for FP32; identical pivot vectors where pivots are well separated.
All norms are Frobenius, accumulated in float64 / complex128.
"""
from __future__ import annotations

import numpy as np

TOL = {np.dtype(np.float32): 1e-5, np.dtype(np.float64): 1e-12, np.dtype(np.complex128): 1e-12}


def tol(dtype):
    return TOL[np.dtype(dtype)]


def _hp(a):
    a = np.asarray(a)
    return a.astype(np.complex128 if np.iscomplexobj(a) else np.float64)


def rel(err, ref):
    d = np.linalg.norm(_hp(ref))
    return float(np.linalg.norm(_hp(err)) / (d if d > 0 else 1.0))


def solve_residual(A, X, B):
    A, X, B = _hp(A), _hp(X), _hp(B)
    den = np.linalg.norm(A) * np.linalg.norm(X)
    return float(np.linalg.norm(A @ X - B) / (den if den > 0 else 1.0))


def lu_residual(A, LU, ipiv):
    """||A - P L U|| / ||A|| from packed getrf output (1-based ipiv)."""
    A, LU = _hp(A), _hp(LU)
    m, n = A.shape
    k = min(m, n)
    L = np.tril(LU[:, :k], -1) + np.eye(m, k)
    U = np.triu(LU[:k, :])
    PA = A.copy()
    for i in range(k):
        p = int(ipiv[i]) - 1
        if p != i:
            PA[[i, p], :] = PA[[p, i], :]
    return rel(PA - L @ U, A)


def chol_residual(A, F, uplo):
    A, F = _hp(A), _hp(F)
    if uplo in ("L", b"L"):
        L = np.tril(F)
        return rel(A - L @ L.conj().T, A)
    U = np.triu(F)
    return rel(A - U.conj().T @ U, A)


def pivots_equal(ipiv, ipiv_ref):
    """(all_equal, first_mismatch_index or -1)."""
    a, b = np.asarray(ipiv), np.asarray(ipiv_ref)
    ne = np.nonzero(a != b)[0]
    return (ne.size == 0, int(ne[0]) if ne.size else -1)
