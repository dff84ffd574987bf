"""Host-side mirror of the reference's `LA::Matrix(T)` API for the hot path.

Crystal is not available in this image, so the reference's host layer (L2-L4 of
SURVEY.md §1) is restated in Python with the same names, argument meaning and error
behaviour, and drives the SAME C ABI through `linalg_b200.ffi.BlasLapackABI`:

  reference (Crystal)                                   here
  -------------------------------------------------------------------------------------
  GeneralMatrix(T)        src/matrix/general_matrix.cr:10    GeneralMatrix
  MatrixFlags             src/matrix/flags.cr:19-34           MatrixFlags
  `lapack` macro          src/linalg/lapack_helper.cr:81-373  _lapack()
  Matrix#* / add_mult / tr_mult!   src/linalg/mult.cr:26-123  __mul__/add_mult/tr_mult_
  solve / det / inv! / solvels / lstsq  src/linalg/linalg.cr:366-560
  lu / lu_factor! / LUMatrix#solve      src/linalg/lu.cr:29-162
  cholesky! / cho_solve                 src/linalg/cholesky.cr:47-90
  qr_raw / qr_r / qr                    src/linalg/qr.cr:34-110

Storage is column-major (numpy order='F'), lda = nrows, exactly the contract that crosses
the FFI in the reference.  Methods ending in `!` in Crystal end in `_` here.
Quirks Q1-Q3 of SURVEY.md §8a (e.g. nrhs = b.nrows in cho_solve) are NOT reproduced where
they would read out of bounds; the well-defined argument values are passed instead.
"""
from __future__ import annotations

import contextlib
import enum
from ctypes import byref, c_char_p, c_int

import numpy as np

from . import ffi

_backend = None


def backend() -> ffi.BlasLapackABI:
    global _backend
    if _backend is None:
        from . import abi
        _backend = abi()
    return _backend


@contextlib.contextmanager
def use_backend(b: ffi.BlasLapackABI):
    """Run the same host code against another library with the same ABI (tests: the oracle)."""
    global _backend
    old = _backend
    _backend = b
    try:
        yield b
    finally:
        _backend = old


class LinAlgError(Exception):
    """src/linalg/linalg.cr:50"""


class MatrixFlags(enum.IntFlag):
    """src/matrix/flags.cr:19-34"""
    NONE = 0
    Symmetric = 1
    Hermitian = 2
    PositiveDefinite = 4
    Orthogonal = 8
    UpperTriangular = 16
    LowerTriangular = 32
    Triangular = 48
    All = 63

    @property
    def symmetric(self): return bool(self & MatrixFlags.Symmetric)
    @property
    def hermitian(self): return bool(self & MatrixFlags.Hermitian)
    @property
    def positive_definite(self): return bool(self & MatrixFlags.PositiveDefinite)
    @property
    def orthogonal(self): return bool(self & MatrixFlags.Orthogonal)
    @property
    def upper_triangular(self): return bool(self & MatrixFlags.UpperTriangular)
    @property
    def lower_triangular(self): return bool(self & MatrixFlags.LowerTriangular)
    @property
    def triangular(self): return self.upper_triangular or self.lower_triangular

    def mult(self, f2):  # flags.cr:66-70
        return self & f2 & (MatrixFlags.UpperTriangular | MatrixFlags.LowerTriangular |
                            MatrixFlags.Symmetric)

    def transpose(self):  # flags.cr:73-90
        r = self
        if self.triangular:
            up, lo = self.upper_triangular, self.lower_triangular
            r = (r | MatrixFlags.LowerTriangular) if up else (r & ~MatrixFlags.LowerTriangular)
            r = (r | MatrixFlags.UpperTriangular) if lo else (r & ~MatrixFlags.UpperTriangular)
        return MatrixFlags(r)


class LSMethod(enum.Enum):
    Auto = 0
    QR = 1
    Orthogonal = 2
    SVD = 3


class LUTranspose(enum.Enum):
    NONE = 0
    Transpose = 1
    ConjTranspose = 2


# lapack_helper.cr:143-183
_WORK_DETECT = {"gels", "geqrf", "getri", "orgqr", "ormqr", "gerqf", "gelqf", "geqlf", "orgrq",
                "orglq", "orgql", "geqp3"}


class MatrixNorm(enum.Enum):      # src/linalg/linalg.cr:703-708
    Frobenius = "F"
    One = "O"
    Inf = "I"
    MaxAbs = "M"


class RankMethod(enum.Enum):
    SVD = 0
    QRP = 1


def _ci(v):
    return byref(c_int(int(v)))


def _lapack(name, letter, *args):
    """The `lapack` macro (lapack_helper.cr:215-372): for WORK_DETECT routines call once with
    lwork = -1, read the size from work[0] (real part), allocate, call again; raise
    LinAlgError when info != 0.  `args` are the routine's arguments without work/lwork/info;
    ints are passed by pointer, bytes as char*, arrays as pointers."""
    fn = backend().lapack(name, letter)
    cargs, keep = [], []
    for a in args:
        if isinstance(a, np.ndarray):
            cargs.append(ffi.ptr(a))
        elif isinstance(a, (bytes, str)):
            b = a if isinstance(a, bytes) else a.encode()
            cargs.append(c_char_p(b))
        else:
            ci = c_int(int(a))
            keep.append(ci)
            cargs.append(byref(ci))
    info = c_int(0)
    extra = []
    if name == "geqp3" and letter == "z":       # rwork, WORK_PARAM1 = 2n (lapack_helper.cr:151)
        rwork = np.zeros(2 * max(1, int(args[1])), dtype=np.float64)
        keep.append(rwork)
        extra = [ffi.ptr(rwork)]
    if name in _WORK_DETECT:
        dt = {"s": np.float32, "d": np.float64, "z": np.complex128}[letter]
        q = np.zeros(1, dtype=dt)
        lw = c_int(-1)
        fn(*cargs, ffi.ptr(q), byref(lw), *extra, byref(info))
        if info.value != 0:
            raise LinAlgError(f"LAPACK.{letter}{name} returned {info.value}")
        lwork = max(1, int(q[0].real))
        work = np.zeros(lwork + 1, dtype=dt)
        guard = dt(0xDEADBEEF)
        work[lwork] = guard  # overrun guard, workpool.cr:37-49
        lw = c_int(lwork)
        fn(*cargs, ffi.ptr(work), byref(lw), *extra, byref(info))
        if work[lwork] != guard:
            raise RuntimeError("worksize guard failed")
    else:
        fn(*cargs, byref(info))
    if info.value != 0:
        raise LinAlgError(f"LAPACK.{letter}{name} returned {info.value}")


class GeneralMatrix:
    """Dense column-major matrix (src/matrix/general_matrix.cr:10-27)."""

    def __init__(self, data, dtype=None, flags=MatrixFlags.NONE):
        a = np.array(data, dtype=dtype, order="F", copy=True)
        if a.ndim == 1:
            a = a.reshape(-1, 1, order="F")
        if a.dtype not in ffi.LETTER:
            a = a.astype(np.complex128 if np.iscomplexobj(a) else np.float64, order="F")
        self.raw = np.asfortranarray(a)
        self.flags = MatrixFlags(flags)

    # -- constructors ---------------------------------------------------------------
    @classmethod
    def wrap(cls, a: np.ndarray, flags=MatrixFlags.NONE):
        m = cls.__new__(cls)
        assert a.flags.f_contiguous and a.dtype in ffi.LETTER
        m.raw = a
        m.flags = MatrixFlags(flags)
        return m

    @classmethod
    def zeros(cls, nrows, ncolumns, dtype=np.float64):
        return cls.wrap(np.zeros((nrows, ncolumns), dtype=dtype, order="F"))

    @classmethod
    def ones(cls, nrows, ncolumns, dtype=np.float64):
        return cls.wrap(np.ones((nrows, ncolumns), dtype=dtype, order="F"))

    @classmethod
    def identity(cls, n, dtype=np.float64):
        return cls.wrap(np.asfortranarray(np.eye(n, dtype=dtype)),
                        MatrixFlags.Symmetric | MatrixFlags.Hermitian | MatrixFlags.Triangular |
                        MatrixFlags.Orthogonal | MatrixFlags.PositiveDefinite)

    @classmethod
    def rand(cls, nrows, ncolumns, rng=None, dtype=np.float64):  # matrix.cr:490-492: U[0,1)
        rng = rng or np.random.default_rng()
        a = rng.random((nrows, ncolumns))
        if np.dtype(dtype) == np.complex128:
            a = a + 1j * rng.random((nrows, ncolumns))
        return cls.wrap(np.asfortranarray(a.astype(dtype)))

    @classmethod
    def columns(cls, cols, dtype=None):
        return cls(np.array(cols, dtype=dtype).T)

    # -- basics ------------------------------------------------------------------------
    @property
    def nrows(self): return self.raw.shape[0]
    @property
    def ncolumns(self): return self.raw.shape[1]
    @property
    def dtype(self): return self.raw.dtype
    @property
    def letter(self): return ffi.LETTER[self.raw.dtype]
    @property
    def square(self): return self.nrows == self.ncolumns
    @property
    def shape(self): return self.raw.shape

    def clone(self):
        return GeneralMatrix.wrap(self.raw.copy(order="F"), self.flags)

    def to_general(self):
        return self.clone()

    def clear_flags(self):
        self.flags = MatrixFlags.NONE

    def assume_(self, flag, value=True):  # flag_checks.cr:5-11
        self.flags = MatrixFlags(self.flags | flag) if value else MatrixFlags(self.flags & ~flag)
        return self

    def __getitem__(self, ij):
        return self.raw[ij]

    def __eq__(self, other):
        o = other.raw if isinstance(other, GeneralMatrix) else np.asarray(other)
        return self.raw.shape == o.shape and bool(np.all(self.raw == o))

    def __repr__(self):
        return f"GeneralMatrix({self.raw!r}, flags={self.flags!r})"

    def t(self):
        return GeneralMatrix.wrap(np.asfortranarray(self.raw.T.copy()), self.flags.transpose())

    transpose = t

    def conjt(self):
        return GeneralMatrix.wrap(np.asfortranarray(self.raw.conj().T.copy()),
                                  self.flags.transpose())

    conjtranspose = conjt

    def tolerance(self):  # matrix.cr:462-471
        amax = float(np.max(np.abs(self.raw))) if self.raw.size else 0.0
        eps = np.finfo(self.raw.real.dtype).eps
        return max(1.0, amax) * self.nrows * self.ncolumns * 10 * eps

    def almost_eq(self, other, eps=None):  # matrix.cr:476-487
        o = other.raw if isinstance(other, GeneralMatrix) else np.asarray(other)
        if self.raw.shape != o.shape:
            return False
        if eps is None:
            ot = other.tolerance() if isinstance(other, GeneralMatrix) else 0.0
            eps = max(self.tolerance(), ot)
        return bool(np.all(np.abs(self.raw - o) <= eps))

    def tril_(self):  # matrix.cr:650-663
        self.raw[:] = np.tril(self.raw)
        self.flags = MatrixFlags.LowerTriangular
        return self

    def triu_(self):
        self.raw[:] = np.triu(self.raw)
        self.flags = MatrixFlags.UpperTriangular
        return self

    def __add__(self, o):
        o = o.raw if isinstance(o, GeneralMatrix) else o
        return GeneralMatrix.wrap(np.asfortranarray(self.raw + o))

    def __sub__(self, o):
        o = o.raw if isinstance(o, GeneralMatrix) else o
        return GeneralMatrix.wrap(np.asfortranarray(self.raw - o))

    # -- flag detection (flag_checks.cr:13-94) --------------------------------------------
    def _check_single(self, flag, eps):
        a = self.raw
        if flag == MatrixFlags.Symmetric:
            return self.square and bool(np.all(np.abs(a - a.T) <= eps))
        if flag == MatrixFlags.Hermitian:
            return self.square and bool(np.all(np.abs(a - a.conj().T) <= eps))
        if flag == MatrixFlags.PositiveDefinite:
            if not (self.square and self._check_single(MatrixFlags.Hermitian, eps)):
                return False
            try:  # flag_checks.cr:32-40: run potrf and catch LinAlgError
                self.clone().cholesky_(dont_clean=True)
                return True
            except LinAlgError:
                return False
        if flag == MatrixFlags.Orthogonal:
            return self.square and (self * self.conjt()).almost_eq(
                GeneralMatrix.identity(self.nrows, self.dtype), eps)
        if flag == MatrixFlags.UpperTriangular:
            return bool(np.all(np.abs(np.tril(a, -1)) <= eps))
        if flag == MatrixFlags.LowerTriangular:
            return bool(np.all(np.abs(np.triu(a, 1)) <= eps))
        return False

    def detect_(self, aflags=MatrixFlags.All, eps=None):
        eps = self.tolerance() if eps is None else eps
        result = True
        for f in (MatrixFlags.Symmetric, MatrixFlags.Hermitian, MatrixFlags.PositiveDefinite,
                  MatrixFlags.Orthogonal, MatrixFlags.LowerTriangular,
                  MatrixFlags.UpperTriangular):
            if aflags & f:
                ok = self._check_single(f, eps)
                self.assume_(f, ok)
                result = result and ok
        return result

    def detect(self, aflags=MatrixFlags.All, eps=None):
        self.detect_(aflags, eps)
        return self

    # -- multiply (mult.cr) ---------------------------------------------------------------
    def add_mult(self, a, b, *, alpha=1.0, beta=1.0):
        """self = alpha*a*b + beta*self via gemm / symm / hemm (mult.cr:26-71)."""
        if a.ncolumns != b.nrows or a.nrows != self.nrows or b.ncolumns != self.ncolumns:
            raise ValueError("matrix size mismatch")
        L = self.letter
        be = backend()
        al, k1 = ffi.scalar_arg(L, alpha)
        bt, k2 = ffi.scalar_arg(L, beta)
        aa, bb = a, b
        herm = L == "z" and (a.flags.hermitian or b.flags.hermitian)
        if herm or a.flags.symmetric or b.flags.symmetric:
            left = a.flags.hermitian if herm else a.flags.symmetric
            side = ffi.LEFT if left else ffi.RIGHT
            if side == ffi.RIGHT:
                aa, bb = bb, aa
            # the reference passes (a.ncolumns, b.nrows) here, correct for square operands
            # only; the true C dimensions are passed instead
            be.symm(L, herm)(ffi.COL_MAJOR, side, ffi.UPPER, self.nrows, self.ncolumns, al,
                             ffi.ptr(aa.raw), aa.nrows, ffi.ptr(bb.raw), bb.nrows, bt,
                             ffi.ptr(self.raw), self.nrows)
        else:
            be.gemm(L)(ffi.COL_MAJOR, ffi.NO_TRANS, ffi.NO_TRANS, a.nrows, b.ncolumns, a.ncolumns,
                       al, ffi.ptr(aa.raw), a.nrows, ffi.ptr(bb.raw), b.nrows, bt,
                       ffi.ptr(self.raw), self.nrows)
        return self

    def tr_mult_(self, a, *, alpha=1.0, left=False):
        """self = alpha*a*self or alpha*self*a via trmm (mult.cr:78-91)."""
        if (self.nrows if left else self.ncolumns) != a.nrows:
            raise ValueError("matrix size should match")
        if not (a.square and a.flags.triangular):
            raise RuntimeError("tr_mult! require square triangular matrix")
        L = self.letter
        al, keep = ffi.scalar_arg(L, alpha)
        uplo = ffi.UPPER if a.flags.upper_triangular else ffi.LOWER
        side = ffi.LEFT if left else ffi.RIGHT
        backend().trmm(L)(ffi.COL_MAJOR, side, uplo, ffi.NO_TRANS, ffi.NON_UNIT, self.nrows,
                          self.ncolumns, al, ffi.ptr(a.raw), a.nrows, ffi.ptr(self.raw),
                          self.nrows)
        return self

    def __mul__(self, m):
        """Matrix#* (mult.cr:106-123); a scalar operand scales elementwise."""
        if not isinstance(m, GeneralMatrix):
            return GeneralMatrix.wrap(np.asfortranarray(self.raw * m))
        if self.ncolumns != m.nrows:
            raise ValueError(f"matrix size should match ({self.shape} * {m.shape})")
        if (self.square and self.flags.triangular) or (m.square and m.flags.triangular):
            if m.square and m.flags.triangular:
                result = self.to_general()
                result.tr_mult_(m, left=False)
            else:
                result = m.to_general()
                result.tr_mult_(self, left=True)
        else:
            result = GeneralMatrix.zeros(self.nrows, m.ncolumns, self.dtype)
            result.add_mult(self, m)
        result.flags = self.flags.mult(m.flags)
        return result

    __matmul__ = __mul__

    def __pow__(self, p):  # power.cr:19-30: repeated squaring
        if not self.square:
            raise ValueError("matrix must be square")
        if p < 0:
            return (self ** (-p)).inv()
        if p > 1 and hasattr(backend().cdll, "la_b200_mat_create"):
            return self._pow_resident(int(p))
        result = GeneralMatrix.identity(self.nrows, self.dtype)
        result.clear_flags()
        base = self
        first = True
        while p > 0:
            if p & 1:
                result = base.clone() if first else result * base
                first = False
            p >>= 1
            if p:
                base = base * base
        return result

    def _pow_resident(self, p):
        """power.cr:19-30 with the operands kept on the device (include/la_b200.h, handles): one
        upload of the base, every product of the repeated squaring through la_b200_?gemm_dev on
        resident matrices (same kernels, same order of products as the pointer path: the result
        is bit-identical), one download of the result."""
        from ctypes import c_double, c_float, c_int64, c_void_p
        L, n = self.letter, self.nrows
        cd = backend().cdll
        cd.la_b200_mat_create.restype = c_void_p
        cd.la_b200_mat_create.argtypes = [c_int, c_int64, c_int64]
        cd.la_b200_mat_ptr.restype, cd.la_b200_mat_ptr.argtypes = c_void_p, [c_void_p]
        cd.la_b200_mat_ld.restype, cd.la_b200_mat_ld.argtypes = c_int64, [c_void_p]
        cd.la_b200_mat_free.restype, cd.la_b200_mat_free.argtypes = None, [c_void_p]
        cd.la_b200_mat_upload.argtypes = [c_void_p, c_void_p, c_int64]
        cd.la_b200_mat_download.argtypes = [c_void_p, c_void_p, c_int64]
        gemm = getattr(cd, f"la_b200_{L}gemm_dev")
        sc = c_void_p if L == "z" else (c_float if L == "s" else c_double)
        gemm.argtypes = [c_int, c_int, c_int64, c_int64, c_int64, sc, c_void_p, c_int64, c_void_p,
                         c_int64, sc, c_void_p, c_int64, c_void_p]
        lacpy = getattr(cd, f"la_b200_{L}lacpy_dev")
        lacpy.argtypes = [c_int, c_int64, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_void_p]
        laset = getattr(cd, f"la_b200_{L}laset_dev")
        laset.argtypes = [c_int, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_int64, c_void_p]
        one, _k1 = ffi.scalar_arg(L, 1.0)
        zbuf = np.zeros(1, dtype=self.dtype)
        hs = [c_void_p(cd.la_b200_mat_create(ord(L), n, n)) for _ in range(4)]
        try:
            if any(h.value is None for h in hs):
                raise MemoryError("la_b200_mat_create failed")
            ld = cd.la_b200_mat_ld(hs[0])
            ptr = cd.la_b200_mat_ptr

            def mult(dst, a, b):           # Matrix#*: beta = 1 on a zeroed result (mult.cr:118-120)
                laset(ord("A"), n, n, ffi.ptr(zbuf), ffi.ptr(zbuf), ptr(dst), ld, None)
                rc = gemm(ffi.NO_TRANS, ffi.NO_TRANS, n, n, n, one, ptr(a), ld, ptr(b), ld, one,
                          ptr(dst), ld, None)
                if rc != 0:
                    raise LinAlgError(f"la_b200_{L}gemm_dev returned {rc}")

            base, result, t1, t2 = hs
            cd.la_b200_mat_upload(base, ffi.ptr(self.raw), n)
            have = False
            while p > 0:
                if p & 1:
                    if not have:           # result = base.clone
                        lacpy(ord("A"), n, n, ptr(base), ld, ptr(result), ld, None)
                        have = True
                    else:
                        mult(t1, result, base)
                        result, t1 = t1, result
                p >>= 1
                if p:
                    mult(t2, base, base)
                    base, t2 = t2, base
            out = GeneralMatrix.zeros(n, n, self.dtype)
            cd.la_b200_mat_download(result, ffi.ptr(out.raw), n)
        finally:
            for h in hs:
                if h.value is not None:
                    cd.la_b200_mat_free(h)
        return out

    # -- LU (lu.cr) --------------------------------------------------------------------------
    def lu(self, *, overwrite_a=False):
        a = self if overwrite_a else self.clone()
        m, n = self.nrows, self.ncolumns
        k = min(m, n)
        ipiv = np.zeros(m, dtype=np.int32)
        _lapack("getrf", self.letter, m, n, a.raw, m, ipiv)
        a.clear_flags()
        piv = list(range(1, m + 1))
        for i in range(k):
            j = ipiv[i] - 1
            piv[i], piv[j] = piv[j], piv[i]
        p = np.zeros((m, m), dtype=self.dtype, order="F")
        for i in range(m):
            p[piv[i] - 1, i] = 1
        l = np.asfortranarray(np.tril(a.raw[:, :k], -1) + np.eye(m, k, dtype=self.dtype))
        u = np.asfortranarray(np.triu(a.raw[:k, :]))
        return GeneralMatrix.wrap(p), GeneralMatrix.wrap(l), GeneralMatrix.wrap(u)

    def lu_factor_(self):
        if not self.square:
            raise ValueError("matrix must be square")
        ipiv = np.zeros(self.nrows, dtype=np.int32)
        _lapack("getrf", self.letter, self.nrows, self.ncolumns, self.raw, self.nrows, ipiv)
        self.clear_flags()
        return LUMatrix(self, ipiv)

    def lu_factor(self):
        return self.clone().lu_factor_()

    def det(self, *, overwrite_a=False):  # linalg.cr:465-475
        if not self.square:
            raise ValueError("matrix must be square")
        if self.flags.triangular:
            return np.prod(np.diag(self.raw))
        lru = self if overwrite_a else self.clone()
        ipiv = np.zeros(self.nrows, dtype=np.int32)
        _lapack("getrf", self.letter, self.nrows, self.nrows, lru.raw, self.nrows, ipiv)
        lru.clear_flags()
        # the reference returns the bare diagonal product (no permutation sign)
        return np.prod(np.diag(lru.raw))

    # -- inverse (linalg.cr:366-397) -------------------------------------------------------------
    def inv_(self):
        if not self.square:
            raise ValueError("can't invert nonsquare matrix")
        n, L = self.nrows, self.letter
        if self.flags.orthogonal:
            self.raw[:] = self.raw.T.copy()
            return self
        if self.flags.triangular:
            uplo = b"U" if self.flags.upper_triangular else b"L"
            _lapack("trtri", L, uplo, b"N", n, self.raw, n)
        elif self.flags.positive_definite:
            uplo = b"U"
            _lapack("potrf", L, uplo, n, self.raw, n)
            _lapack("potri", L, uplo, n, self.raw, n)
            # adjust_symmetric: mirror the computed triangle
            iu = np.triu_indices(n, 1)
            self.raw.T[iu] = self.raw[iu].conj()
        else:
            ipiv = np.zeros(n, dtype=np.int32)
            _lapack("getrf", L, n, n, self.raw, n, ipiv)
            _lapack("getri", L, n, self.raw, n, ipiv)
        return self

    def inv(self):
        return self.clone().inv_()

    # -- solve (linalg.cr:425-450) ---------------------------------------------------------------
    def solve(self, b, *, overwrite_a=False, overwrite_b=False):
        if self.nrows != b.nrows:
            raise ValueError("nrows of a and b must match")
        if not self.square:
            raise ValueError("a must be square")
        a = self if overwrite_a else self.clone()
        x = b if overwrite_b else b.clone()
        n, L = self.nrows, self.letter
        if self.flags.triangular:
            uplo = b"U" if self.flags.upper_triangular else b"L"
            _lapack("trtrs", L, uplo, b"N", b"N", n, b.ncolumns, a.raw, n, x.raw, b.nrows)
        elif self.flags.positive_definite:
            _lapack("posv", L, b"U", n, b.ncolumns, a.raw, n, x.raw, b.nrows)
        else:
            ipiv = np.zeros(n, dtype=np.int32)
            _lapack("gesv", L, n, b.ncolumns, a.raw, n, ipiv, x.raw, b.nrows)
        a.clear_flags()
        x.clear_flags()
        return x

    # -- Cholesky (cholesky.cr) ----------------------------------------------------------------------
    def cholesky_(self, *, lower=False, dont_clean=False):
        if not self.square:
            raise ValueError("Matrix must be square for cholesky decomposition")
        _lapack("potrf", self.letter, b"L" if lower else b"U", self.nrows, self.raw, self.nrows)
        if lower:
            if dont_clean:
                self.flags = MatrixFlags.LowerTriangular
            else:
                self.tril_()
        else:
            if dont_clean:
                self.flags = MatrixFlags.UpperTriangular
            else:
                self.triu_()
        return self

    def cholesky(self, *, lower=False, dont_clean=False):
        m = self.clone().cholesky_(lower=lower, dont_clean=dont_clean)
        self.assume_(MatrixFlags.PositiveDefinite)
        return m

    def cho_solve(self, b, *, overwrite_b=False):
        if self.nrows != b.nrows:
            raise ValueError("nrows of a and b must match")
        if not self.square:
            raise ValueError("a must be square")
        if not self.flags.triangular:
            raise ValueError("a must be triangular")
        x = b if overwrite_b else b.clone()
        n = self.nrows
        uplo = b"U" if self.flags.upper_triangular else b"L"
        _lapack("potrs", self.letter, uplo, n, b.ncolumns, self.raw, n, x.raw, b.nrows)
        return x

    # -- QR (qr.cr) --------------------------------------------------------------------------------------
    def qr_raw(self, *, overwrite_a=False, pivoting=False):   # qr.cr:34-54
        a = self if overwrite_a else self.clone()
        m, n = a.nrows, a.ncolumns
        tau = np.zeros(n, dtype=self.dtype)
        if pivoting:
            jpvt = np.zeros(n, dtype=np.int32)
            _lapack("geqp3", self.letter, m, n, a.raw, m, jpvt, tau)
            a.clear_flags()
            return a, tau, jpvt
        _lapack("geqrf", self.letter, m, n, a.raw, m, tau)
        a.clear_flags()
        return a, tau

    def qr_r(self, *, overwrite_a=False, pivoting=False):
        if pivoting:
            a, _, pvt = self.qr_raw(overwrite_a=overwrite_a, pivoting=True)
            a.triu_()
            return a, pvt
        a, _ = self.qr_raw(overwrite_a=overwrite_a)
        a.triu_()
        return a

    # -- RQ, LQ, QL (rq_lq_ql.cr:3-39) ------------------------------------------------------------
    def _decomposition(self, letters, overwrite_a):
        a = self if overwrite_a else self.clone()
        m, n = a.nrows, a.ncolumns
        tau = np.zeros(n, dtype=self.dtype)
        _lapack("ge" + letters + "f", self.letter, m, n, a.raw, m, tau)
        a.clear_flags()
        r = a.clone()
        r.tril_() if "l" in letters else r.triu_()
        k = min(m, n)
        _lapack("org" + letters, self.letter, m, n, k, a.raw, m, tau)
        a.assume_(MatrixFlags.Orthogonal)
        return (a, r) if letters[0] == "q" else (r, a)

    def rq(self, *, overwrite_a=False): return self._decomposition("rq", overwrite_a)

    def lq(self, *, overwrite_a=False): return self._decomposition("lq", overwrite_a)

    def ql(self, *, overwrite_a=False): return self._decomposition("ql", overwrite_a)

    # -- norms and rank (linalg.cr:721-780) -----------------------------------------------------------
    def norm(self, kind=MatrixNorm.Frobenius):
        let, L = kind.value.encode(), self.letter
        n = self.nrows
        work = np.zeros(max(1, n), dtype=np.float32 if L == "s" else np.float64)
        ab = backend()
        if self.flags.triangular:
            uplo = b"U" if self.flags.upper_triangular else b"L"
            return float(ab.norm("lantr", L)(let, uplo, b"N", _ci(self.nrows), _ci(self.ncolumns),
                                             ffi.ptr(self.raw), _ci(self.nrows), ffi.ptr(work)))
        if L == "z" and self.flags.hermitian:
            return float(ab.norm("lanhe", L)(let, b"U", _ci(n), ffi.ptr(self.raw), _ci(n), ffi.ptr(work)))
        if self.flags.symmetric:
            return float(ab.norm("lansy", L)(let, b"U", _ci(n), ffi.ptr(self.raw), _ci(n), ffi.ptr(work)))
        return float(ab.norm("lange", L)(let, _ci(self.nrows), _ci(self.ncolumns), ffi.ptr(self.raw),
                                         _ci(self.nrows), ffi.ptr(work)))

    def rank(self, eps=None, *, method=RankMethod.QRP, overwrite_a=False):
        eps = self.tolerance() if eps is None else eps
        if method is not RankMethod.QRP:
            raise NotImplementedError("svdvals (gesdd) is outside the hot path (SURVEY.md §2)")
        a, _ = self.qr_r(overwrite_a=overwrite_a, pivoting=True)
        return int(np.count_nonzero(np.abs(np.diagonal(a.raw)) > eps))

    def qr(self, *, overwrite_a=False):
        a, tau = self.qr_raw(overwrite_a=overwrite_a)
        r = a.clone()
        r.triu_()
        m, n = a.nrows, a.ncolumns
        k = min(m, n)
        if n > m:  # orgqr needs n <= m: Q is m x m, built from the leading m columns
            q = GeneralMatrix.wrap(np.asfortranarray(a.raw[:, :m].copy()))
            _lapack("orgqr", self.letter, m, m, k, q.raw, m, tau)
        else:
            q = a
            _lapack("orgqr", self.letter, m, n, k, q.raw, m, tau)
        q.assume_(MatrixFlags.Orthogonal)
        return q, r

    # -- least squares (linalg.cr:493-560) -------------------------------------------------------------------
    def solvels(self, b, *, overwrite_a=False, overwrite_b=False):
        if self.nrows != b.nrows:
            raise ValueError("nrows of a and b must match")
        a = self if overwrite_a else self.clone()
        if self.ncolumns > self.nrows:  # make room for the solution
            xr = np.zeros((self.ncolumns, b.ncolumns), dtype=self.dtype, order="F")
            xr[: self.nrows, :] = b.raw
            x = GeneralMatrix.wrap(xr)
        else:
            x = b if overwrite_b else b.clone()
        _lapack("gels", self.letter, b"N", self.nrows, self.ncolumns, b.ncolumns, a.raw,
                self.nrows, x.raw, x.nrows)
        a.clear_flags()
        x.clear_flags()
        return x

    def lstsq(self, b, method=LSMethod.Auto, **kw):
        if method in (LSMethod.Auto, LSMethod.QR):
            return self.solvels(b, **kw), 0, []
        raise NotImplementedError("gelsd/gelsy are outside the hot path (SURVEY.md §2)")


class LUMatrix:
    """src/linalg/lu.cr:92-162"""

    def __init__(self, a: GeneralMatrix, ipiv: np.ndarray):
        self.a = a
        self.ipiv = ipiv

    @property
    def size(self):
        return self.a.nrows

    def solve(self, b, transpose=LUTranspose.NONE, *, overwrite_b=False):
        if self.a.nrows != b.nrows:
            raise ValueError("nrows of a and b must match")
        trans = {LUTranspose.NONE: b"N", LUTranspose.Transpose: b"T",
                 LUTranspose.ConjTranspose: b"C"}[transpose]
        x = b if overwrite_b else b.clone()
        _lapack("getrs", self.a.letter, trans, self.size, b.ncolumns, self.a.raw, self.size,
                self.ipiv, x.raw, x.nrows)
        x.clear_flags()
        return x


# module-level wrappers (linalg.cr:57-139)
def lapack_version():
    return backend().ilaver()


def solve(a, b, **kw): return a.solve(b, **kw)
def inv(a): return a.inv()
def solvels(a, b, **kw): return a.solvels(b, **kw)
def lstsq(a, b, method=LSMethod.Auto, **kw): return a.lstsq(b, method, **kw)
def cho_solve(a, b, **kw): return a.cho_solve(b, **kw)


def GMat(data): return GeneralMatrix(data, dtype=np.float64)
def GMat32(data): return GeneralMatrix(data, dtype=np.float32)
def GMatComplex(data): return GeneralMatrix(data, dtype=np.complex128)
