"""TEST INFRASTRUCTURE — CPU restatement of the reference's matrix functions.

numpy restatement of konovod/linalg's scaling-and-squaring matrix exponential
(`src/linalg/expm.cr`, itself a conversion of Al-Mohy & Higham's EXPM_NEW, SIAM J. Matrix Anal.
Appl. 31(3), 2009) and of the functions built on it (`src/linalg/matfun.cr`).  Only `tests/` may
import this module: it is the checker for `linalg_b200.la.GeneralMatrix.expm` and friends, never
part of the product path.

Pinned by `tests/test_expm.py` against every golden vector of the reference's own specs
(`spec/expm_spec.cr`, `spec/matfun_spec.cr`) and against `scipy.linalg.expm` (the same
algorithm family) on random matrices.

What is restated: the general (full matrix, no Schur form) path, the zero / diagonal / 2x2
special cases.  Not restated: `schur_fact: true` (needs the eigenvalue routines, out of the
hot-path scope) and the triangular squaring phase `expm_sqtri` (`expm.cr:232-294`).
"""
from __future__ import annotations

import math

import numpy as np

# expm.cr:4-26
COEFF = (1.0 / 100800, 1.0 / 10059033600, 1.0 / 4487938430976000,
         1.0 / 5914384781877411840000.0, 1.0 / 113250775606021113483283660800000000.0)
M_VALS = (3, 5, 7, 9, 13)
THETA = (1.495585217958292e-002, 2.539398330063230e-001, 9.504178996162932e-001,
         2.097847961257068e+000, 4.250000000000000e+000)

# expm.cr:168-191
PADE = {
    3: (120.0, 60.0, 12.0, 1.0),
    5: (30240.0, 15120.0, 3360.0, 420.0, 30.0, 1.0),
    7: (17297280.0, 8648640.0, 1995840.0, 277200.0, 25200.0, 1512.0, 56.0, 1.0),
    9: (17643225600.0, 8821612800.0, 2075673600.0, 302702400.0, 30270240.0,
        2162160.0, 110880.0, 3960.0, 90.0, 1.0),
    13: (64764752532480000.0, 32382376266240000.0, 7771770303897600.0,
         1187353796428800.0, 129060195264000.0, 10559470521600.0,
         670442572800.0, 33522128640.0, 1323241920.0,
         40840800.0, 960960.0, 16380.0, 182.0, 1.0),
}


def norm1(a):
    return float(np.abs(a).sum(axis=0).max()) if a.size else 0.0


def norm_am(a, m):
    """expm.cr:193-217 — ||A^m||_1 through m products with a vector of ones.  A matrix with an
    entry whose real part is negative is replaced by its modulus first; note that the reference's
    test is on the REAL part only, so a complex matrix with non-negative real parts is used as
    it is (the estimate is then the modulus of complex sums) — restated as written."""
    a = np.asarray(a)
    b = a if bool(np.all(a.real >= 0)) else np.abs(a)
    e = np.ones((a.shape[0], 1), dtype=b.dtype)
    bt = b.T
    for _ in range(m):
        e = bt @ e
    return float(np.abs(e).max())


def eval_alpha(a, k, eps):
    """expm.cr:112-117"""
    u = eps / 2
    alpha = COEFF[k - 1] * norm_am(np.abs(a), 2 * M_VALS[k - 1] + 1) / norm1(a)
    if alpha == 0.0:
        return 0            # log2(0) = -inf in the reference: {ceil(-inf), 0}.max = 0
    return max(int(math.ceil(math.log2(alpha / u) / (2 * M_VALS[k - 1]))), 0)


def pade(a, m, a2, a4=None, a6=None):
    """expm.cr:119-166"""
    c = PADE[m]
    n = a.shape[0]
    eye = np.eye(n, dtype=a.dtype)
    if m in (3, 5, 7, 9):
        m2 = (m + 3) // 2
        pw = [eye, a2]
        if a4 is not None:
            pw.append(a4)
        if a6 is not None:
            pw.append(a6)
        while len(pw) < m2:
            pw.append(pw[-1] @ a2)
        u = np.zeros_like(a)
        v = np.zeros_like(a)
        for j in range(m + 1, 1, -2):
            u = u + c[j - 1] * pw[j // 2 - 1]
        u = a @ u
        for j in range(m, 0, -2):
            v = v + c[j - 1] * pw[(j + 1) // 2 - 1]
    else:
        u = a @ (a6 @ (c[13] * a6 + c[11] * a4 + c[9] * a2) + c[7] * a6 + c[5] * a4 + c[3] * a2 + c[1] * eye)
        v = a6 @ (c[12] * a6 + c[10] * a4 + c[8] * a2) + c[6] * a6 + c[4] * a4 + c[2] * a2 + c[0] * eye
    return np.linalg.inv(v - u) @ (u + v)


def _sinch(x):
    return 1.0 if x == 0 else np.sinh(x) / x


def expm_2by2_full(a):
    """expm.cr:315-329"""
    p, b, c, d = a[0, 0], a[0, 1], a[1, 0], a[1, 1]
    delta = np.sqrt((p - d) * (p - d) + 4 * b * c + 0j) if np.iscomplexobj(a) else None
    if delta is None:
        disc = (p - d) * (p - d) + 4 * b * c
        delta = np.sqrt(disc) if disc >= 0 else np.sqrt(disc + 0j)
    k = np.exp((p + d) / 2)
    out = np.array([[k * (np.cosh(delta / 2) + (p - d) / 2 * _sinch(delta / 2)), k * b * _sinch(delta / 2)],
                    [k * c * _sinch(delta / 2), k * (np.cosh(delta / 2) + (d - p) / 2 * _sinch(delta / 2))]])
    return out if np.iscomplexobj(a) else out.real.astype(a.dtype)


def expm(a, diagonal=False):
    """expm.cr:37-110, general path.  `diagonal`: the matrix carries the reference's Diagonal flag."""
    a = np.array(a)
    if not np.issubdtype(a.dtype, np.inexact):
        a = a.astype(np.float64)
    n = a.shape[0]
    assert a.shape == (n, n)
    eps = float(np.finfo(a.dtype).eps)
    if norm1(a) == 0:
        return np.eye(n, dtype=a.dtype)
    if diagonal:
        return np.diag(np.exp(np.diagonal(a))).astype(a.dtype)
    if n == 2:
        return expm_2by2_full(a).astype(a.dtype)
    a2 = a @ a
    eta1 = max(norm_am(a2, 2) ** (1.0 / 4), norm_am(a2, 3) ** (1.0 / 6))
    if eta1 <= THETA[0] and eval_alpha(a, 1, eps) == 0:
        return pade(a, 3, a2)
    a4 = a2 @ a2
    eta2 = max(norm_am(a4, 1) ** (1.0 / 4), norm_am(a2, 3) ** (1.0 / 6))
    if eta2 <= THETA[1] and eval_alpha(a, 2, eps) == 0:
        return pade(a, 5, a2, a4)
    a6 = a2 @ a4
    eta3 = max(norm_am(a6, 1) ** (1.0 / 6), norm_am(a4, 2) ** (1.0 / 8))
    for i in (2, 3):
        if eta3 <= THETA[i] and eval_alpha(a, i + 1, eps) == 0:
            return pade(a, M_VALS[i], a2, a4, a6)
    eta4 = max(norm_am(a4, 2) ** (1.0 / 8), norm_am(a2, 5) ** (1.0 / 10))
    eta5 = min(eta3, eta4)
    s = max(int(math.ceil(math.log2(eta5 / THETA[-1]))), 0) if eta5 > 0 else 0
    s += eval_alpha(a / 2.0 ** s, 5, eps)
    a, a2, a4, a6 = a / 2.0 ** s, a2 / 2.0 ** (2 * s), a4 / 2.0 ** (4 * s), a6 / 2.0 ** (6 * s)
    f = pade(a, 13, a2, a4, a6)
    for _ in range(s):
        f = f @ f
    return f


# matfun.cr:20-56
def cosm(a):
    a = np.asarray(a)
    if np.iscomplexobj(a):
        return 0.5 * (expm(1j * a) + expm(-1j * a))
    return expm(1j * a.astype(np.complex128)).real


def sinm(a):
    a = np.asarray(a)
    if np.iscomplexobj(a):
        return -0.5j * (expm(1j * a) - expm(-1j * a))
    return expm(1j * a.astype(np.complex128)).imag


def tanm(a):
    return sinm(a) @ np.linalg.inv(cosm(a))


def coshm(a):
    a = np.asarray(a)
    return 0.5 * (expm(a) + expm(-a))


def sinhm(a):
    a = np.asarray(a)
    return 0.5 * (expm(a) - expm(-a))


def tanhm(a):
    return sinhm(a) @ np.linalg.inv(coshm(a))
