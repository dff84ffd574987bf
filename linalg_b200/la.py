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



class _DeviceChain:
    """Resident operands for a chain of products, element-wise steps and one solve (SURVEY.md §8 f3;
    include/la_b200.h Part 3 + handles): square n x n matrices and n x 1 vectors of one element
    type live in device handles, every step is one `la_b200_*_dev` call on the default stream,
    only scalars and the final result cross PCIe."""

    def __init__(self, letter, n, dtype):
        from ctypes import c_double, c_float, c_int64, c_void_p
        self.L, self.n, self.dtype = letter, n, np.dtype(dtype)
        cd = self.cd = backend().cdll
        if not hasattr(cd, "la_b200_mat_create"):
            raise RuntimeError("the backend has no device-matrix handles")
        cd.la_b200_mat_create.restype = c_void_p
        cd.la_b200_mat_create.argtypes = [c_int, c_int64, c_int64]
        cd.la_b200_mat_ptr.restype, cd.la_b200_mat_ptr.argtypes = c_void_p, [c_void_p]
        cd.la_b200_mat_ld.restype, cd.la_b200_mat_ld.argtypes = c_int64, [c_void_p]
        cd.la_b200_mat_free.restype, cd.la_b200_mat_free.argtypes = None, [c_void_p]
        cd.la_b200_mat_upload.argtypes = [c_void_p, c_void_p, c_int64]
        cd.la_b200_mat_download.argtypes = [c_void_p, c_void_p, c_int64]
        sc = c_void_p if letter == "z" else (c_float if letter == "s" else c_double)
        f = lambda name: getattr(cd, f"la_b200_{letter}{name}_dev")
        self._gemm = f("gemm")
        self._gemm.argtypes = [c_int, c_int, c_int64, c_int64, c_int64, sc, c_void_p, c_int64,
                               c_void_p, c_int64, sc, c_void_p, c_int64, c_void_p]
        self._geadd = f("geadd")
        self._geadd.argtypes = [c_int64, c_int64, c_void_p, c_void_p, c_int64, c_void_p, c_void_p,
                                c_int64, c_void_p]
        self._geabs = f("geabs")
        self._geabs.argtypes = [c_int64, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_void_p]
        self._laset = f("laset")
        self._laset.argtypes = [c_int, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_int64, c_void_p]
        self._lange = f("lange")
        self._lange.argtypes = [c_int, c_int64, c_int64, c_void_p, c_int64, c_void_p, c_void_p]
        self._getrf = f("getrf")
        self._getrf.argtypes = [c_int64, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_void_p]
        self._getrs = f("getrs")
        self._getrs.argtypes = [c_int, c_int64, c_int64, c_void_p, c_int64, c_void_p, c_void_p,
                                c_int64, c_void_p]
        self._handles = []
        self._one, self._k1 = ffi.scalar_arg(letter, 1.0)
        self._zero, self._k0 = ffi.scalar_arg(letter, 0.0)
        self._scalar = self._handle("d", 1, 1)          # a device double for norms
        self._ipiv = self._handle("s", n, 1)            # 4-byte slots: the pivots (int32)
        self._info = self._handle("s", 1, 1)

    def _handle(self, letter, rows, cols):
        from ctypes import c_void_p
        h = c_void_p(self.cd.la_b200_mat_create(ord(letter), rows, cols))
        if h.value is None:
            self.close()
            raise MemoryError("la_b200_mat_create failed")
        self._handles.append(h)
        return h

    def close(self):
        for h in self._handles:
            self.cd.la_b200_mat_free(h)
        self._handles = []

    def __enter__(self): return self
    def __exit__(self, *exc): self.close()

    def matrix(self): return self._handle(self.L, self.n, self.n)
    def vector(self): return self._handle(self.L, self.n, 1)
    def _p(self, h): return self.cd.la_b200_mat_ptr(h)
    def _ld(self, h): return self.cd.la_b200_mat_ld(h)

    def _check(self, rc, what):
        if rc != 0:
            raise LinAlgError(f"la_b200_{self.L}{what}_dev returned {rc}")

    def upload(self, h, a):
        a = np.asfortranarray(a, dtype=self.dtype)
        self._check(self.cd.la_b200_mat_upload(h, ffi.ptr(a), a.shape[0]), "mat_upload")

    def download(self, h, rows, cols):
        out = np.zeros((rows, cols), dtype=self.dtype, order="F")
        self._check(self.cd.la_b200_mat_download(h, ffi.ptr(out), rows), "mat_download")
        return out

    def _sc(self, v):
        buf = np.array([v], dtype=self.dtype)
        return buf

    def fill(self, h, offdiag, diag, cols=None):
        a, b = self._sc(offdiag), self._sc(diag)
        cols = self.n if cols is None else cols
        self._check(self._laset(ord("A"), self.n, cols, ffi.ptr(a), ffi.ptr(b), self._p(h),
                                self._ld(h), None), "laset")

    def mul(self, dst, a, b, cols=None, trans_a=False):
        """dst <- op(a) b: Matrix#* is beta = 1 on a zeroed result (mult.cr:118-120)."""
        cols = self.n if cols is None else cols
        self.fill(dst, 0, 0, cols)
        self._check(self._gemm(ffi.TRANS if trans_a else ffi.NO_TRANS, ffi.NO_TRANS, self.n, cols,
                               self.n, self._one, self._p(a), self._ld(a), self._p(b), self._ld(b),
                               self._one, self._p(dst), self._ld(dst), None), "gemm")

    def axpby(self, dst, alpha, a, beta=0.0):
        """dst <- alpha a + beta dst"""
        al, be = self._sc(alpha), self._sc(beta)
        self._check(self._geadd(self.n, self.n, ffi.ptr(al), self._p(a), self._ld(a), ffi.ptr(be),
                                self._p(dst), self._ld(dst), None), "geadd")

    def absolute(self, dst, a):
        self._check(self._geabs(self.n, self.n, self._p(a), self._ld(a), self._p(dst), self._ld(dst),
                                None), "geabs")

    def norm(self, h, kind, cols=None):
        from ctypes import c_double
        cols = self.n if cols is None else cols
        self._check(self._lange(ord(kind), self.n, cols, self._p(h), self._ld(h),
                                self._p(self._scalar), None), "lange")
        out = np.zeros(1, dtype=np.float64)
        self._check(self.cd.la_b200_mat_download(self._scalar, ffi.ptr(out), 1), "mat_download")
        return float(out[0])

    def solve(self, a, b):
        """b <- a^-1 b through ?getrf / ?getrs on the resident matrices (a is overwritten)."""
        n = self.n
        self._check(self._getrf(n, n, self._p(a), self._ld(a), self._p(self._ipiv),
                                self._p(self._info), None), "getrf")
        info = np.zeros(1, dtype=np.float32)
        self._check(self.cd.la_b200_mat_download(self._info, ffi.ptr(info), 1), "mat_download")
        if int(info.view(np.int32)[0]) != 0:
            raise LinAlgError(f"?getrf returned {int(info.view(np.int32)[0])}")
        self._check(self._getrs(ffi.NO_TRANS, n, n, self._p(a), self._ld(a), self._p(self._ipiv),
                                self._p(b), self._ld(b), None), "getrs")


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

    def _clear_triangle(self, upper):
        """Zero the strictly upper / lower triangle in place through ?laset_ (the
        reference's `tril!` / `triu!` are scalar-callback loops, matrix.cr:650-663): ?laset on
        the sub-matrix shifted by one column / row sets exactly the strict triangle."""
        m, n = self.raw.shape
        if m == 0 or n == 0 or (upper and n < 2) or (not upper and m < 2):
            return
        zero = np.zeros(1, dtype=self.dtype)
        fn = backend().lapack("laset", self.letter)
        item = self.raw.itemsize
        base = self.raw.ctypes.data
        if upper:   # strictly upper part of A = upper triangle (with diagonal) of A[:, 1:]
            sub = base + m * item
            fn(b"U", _ci(min(m, n - 1)), _ci(n - 1), ffi.ptr(zero), ffi.ptr(zero), sub, _ci(m))
        else:       # strictly lower part of A = lower triangle (with diagonal) of A[1:, :]
            sub = base + item
            fn(b"L", _ci(m - 1), _ci(min(n, m - 1)), ffi.ptr(zero), ffi.ptr(zero), sub, _ci(m))

    def tril_(self):  # matrix.cr:650-663
        self._clear_triangle(upper=True)
        self.flags = MatrixFlags.LowerTriangular
        return self

    def triu_(self):
        self._clear_triangle(upper=False)
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

    # -- matrix functions (expm.cr, matfun.cr) ------------------------------------------------------
    # expm.cr:4-26, 168-191
    _EXPM_COEFF = (1.0 / 100800, 1.0 / 10059033600, 1.0 / 4487938430976000,
                   1.0 / 5914384781877411840000.0, 1.0 / 113250775606021113483283660800000000.0)
    _EXPM_M = (3, 5, 7, 9, 13)
    _EXPM_THETA = (1.495585217958292e-002, 2.539398330063230e-001, 9.504178996162932e-001,
                   2.097847961257068e+000, 4.250000000000000e+000)
    _EXPM_PADE = {
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
    EXPM_RESIDENT_FROM = 64     # orders from which expm keeps its operands on the device

    def expm(self, *, schur_fact=False):
        """expm.cr:37-110 — scaling and squaring with Pade approximants (Al-Mohy & Higham 2009),
        same degree selection, same 1-norm estimates of powers (`normAm`), same special cases
        (zero matrix, Diagonal flag, 2 x 2 closed forms).  From order EXPM_RESIDENT_FROM on the
        whole chain — products, |A|, linear combinations, the final solve and the squarings — runs
        on device-resident operands (one upload, one download).  Differences from the reference,
        all at rounding level: (V - U)^-1 (U + V) is a ?getrf / ?getrs solve, not an explicit
        inverse followed by a product; a matrix flagged (one-sided) triangular takes this general
        path instead of `expm_sqtri` (expm.cr:232-294); `schur_fact: true` needs the eigenvalue
        routines, which are outside the hot path, and is refused."""
        if not self.square:
            raise ValueError("Matrix must be square for expm")
        if schur_fact:
            raise NotImplementedError("expm(schur_fact: true) needs ?gees (outside the hot path)")
        n = self.nrows
        resident = n >= self.EXPM_RESIDENT_FROM and hasattr(backend().cdll, "la_b200_mat_create")
        # norm(One) == 0 (the resident chain takes the norm on the device, after its upload)
        if n == 0 or (not resident and float(np.abs(self.raw).sum(axis=0).max()) == 0.0):
            return GeneralMatrix.identity(n, self.dtype)
        if self.flags.upper_triangular and self.flags.lower_triangular:    # flags.diagonal?
            return GeneralMatrix(np.diag(np.exp(np.diagonal(self.raw))).astype(self.dtype))
        if n == 2:
            return GeneralMatrix(self._expm_2by2())
        if resident:
            return GeneralMatrix(self._expm_resident())
        return GeneralMatrix(self._expm_host())

    def _expm_2by2(self):       # expm.cr:297-357
        a, b, c, d = (self.raw[0, 0], self.raw[0, 1], self.raw[1, 0], self.raw[1, 1])
        sinch = lambda x: 1.0 if x == 0 else np.sinh(x) / x
        if self.flags.upper_triangular:                                    # expmT2by2!
            ave = ((a + d) / 2).real
            df = abs(a - d) / 2
            if max(ave, df) < np.log(np.finfo(self.raw.real.dtype).max):
                x12 = b * np.exp((a + d) / 2) * sinch((d - a) / 2)
            else:
                x12 = b * (np.exp(d) - np.exp(a)) / (d - a)
            return np.array([[np.exp(a), x12], [0.0, np.exp(d)]], dtype=self.dtype)
        disc = (a - d) * (a - d) + 4 * b * c
        delta = np.sqrt(disc + 0j) if (np.iscomplexobj(self.raw) or disc < 0) else np.sqrt(disc)
        k = np.exp((a + d) / 2)
        out = np.array([[k * (np.cosh(delta / 2) + (a - d) / 2 * sinch(delta / 2)), k * b * sinch(delta / 2)],
                        [k * c * sinch(delta / 2), k * (np.cosh(delta / 2) + (d - a) / 2 * sinch(delta / 2))]])
        return (out if np.iscomplexobj(self.raw) else out.real).astype(self.dtype)

    @staticmethod
    def _expm_alpha_steps(alpha, eps, m):      # expm.cr:112-117
        import math
        if alpha == 0.0:
            return 0
        return max(int(math.ceil(math.log2(alpha / (eps / 2)) / (2 * m))), 0)

    def _expm_plan(self, norm_am, norm_am_abs_a, norm1_a):
        """The degree / scaling decisions of expm.cr:58-98, written once for both paths.
        `norm_am(which, m)` estimates ||X^m||_1 for X in {"a2", "a4", "a6"} (computing the power
        on demand), `norm_am_abs_a(m, scale)` does it for |A| / scale.  Returns (degree, s)."""
        import math
        eps = float(np.finfo(self.raw.real.dtype).eps)
        TH, MV, CO = self._EXPM_THETA, self._EXPM_M, self._EXPM_COEFF
        steps = lambda k, scale=1.0: self._expm_alpha_steps(
            CO[k - 1] * norm_am_abs_a(2 * MV[k - 1] + 1, scale) / (norm1_a / scale), eps, MV[k - 1])
        eta1 = max(norm_am("a2", 2) ** (1.0 / 4), norm_am("a2", 3) ** (1.0 / 6))
        if eta1 <= TH[0] and steps(1) == 0:
            return 3, 0
        eta2 = max(norm_am("a4", 1) ** (1.0 / 4), norm_am("a2", 3) ** (1.0 / 6))
        if eta2 <= TH[1] and steps(2) == 0:
            return 5, 0
        eta3 = max(norm_am("a6", 1) ** (1.0 / 6), norm_am("a4", 2) ** (1.0 / 8))
        for i in (2, 3):
            if eta3 <= TH[i] and steps(i + 1) == 0:
                return MV[i], 0
        eta4 = max(norm_am("a4", 2) ** (1.0 / 8), norm_am("a2", 5) ** (1.0 / 10))
        eta5 = min(eta3, eta4)
        s = max(int(math.ceil(math.log2(eta5 / TH[-1]))), 0) if eta5 > 0 else 0
        s += steps(5, 2.0 ** s)
        return 13, s

    def _expm_host(self):
        """The chain with every product through Matrix#* (cblas_?gemm on host pointers), the
        element-wise steps in host memory as in the reference."""
        A = self.clone()
        A.clear_flags()
        n, dt = self.nrows, self.dtype
        pw = {}

        def power(which):
            if which not in pw:
                if which == "a2":
                    pw[which] = A * A
                elif which == "a4":
                    pw[which] = power("a2") * power("a2")
                else:
                    pw[which] = power("a2") * power("a4")
            return pw[which]

        def norm_am_of(mat, m):          # expm.cr:193-217 (test on the real parts, as written)
            b = mat if bool(np.all(mat.real >= 0)) else np.abs(mat)
            e = GeneralMatrix.ones(n, 1, b.dtype)
            bt = GeneralMatrix(np.asfortranarray(b.T))
            for _ in range(m):
                e = bt * e
            return float(np.abs(e.raw).max())

        absa = np.abs(A.raw)
        norm1_a = float(absa.sum(axis=0).max())
        deg, s = self._expm_plan(lambda w, m: norm_am_of(power(w).raw, m),
                                 lambda m, scale: norm_am_of(absa / scale, m), norm1_a)
        c = self._EXPM_PADE[deg]
        eye = np.eye(n, dtype=dt)
        a = A.raw / 2.0 ** s
        a2 = power("a2").raw / 2.0 ** (2 * s)
        G = lambda x: GeneralMatrix(np.asfortranarray(x.astype(dt)))
        if deg == 13:
            a4, a6 = power("a4").raw / 2.0 ** (4 * s), power("a6").raw / 2.0 ** (6 * s)
            u = (G(a) * (G(a6) * G(c[13] * a6 + c[11] * a4 + c[9] * a2) +
                         G(c[7] * a6 + c[5] * a4 + c[3] * a2 + c[1] * eye))).raw
            v = (G(a6) * G(c[12] * a6 + c[10] * a4 + c[8] * a2)).raw + \
                c[6] * a6 + c[4] * a4 + c[2] * a2 + c[0] * eye
        else:
            m2 = (deg + 3) // 2
            pws = [eye, a2]
            if deg >= 5:
                pws.append(power("a4").raw)
            if deg >= 7:
                pws.append(power("a6").raw)
            while len(pws) < m2:
                pws.append((G(pws[-1]) * G(a2)).raw)
            u = np.zeros((n, n), dtype=dt)
            v = np.zeros((n, n), dtype=dt)
            for j in range(deg + 1, 1, -2):
                u = u + c[j - 1] * pws[j // 2 - 1]
            u = (G(a) * G(u)).raw
            for j in range(deg, 0, -2):
                v = v + c[j - 1] * pws[(j + 1) // 2 - 1]
        f = G(v - u).solve(G(u + v))
        for _ in range(s):
            f = f * f
        return f.raw

    def _expm_resident(self):
        """The same chain on device-resident operands (`_DeviceChain`)."""
        n, dt, L = self.nrows, self.dtype, self.letter
        cplx = np.iscomplexobj(self.raw)
        with _DeviceChain(L, n, dt) as dc:
            A, absA, t1, t2, t3 = (dc.matrix() for _ in range(5))
            e1, e2 = dc.vector(), dc.vector()
            dc.upload(A, self.raw)
            dc.absolute(absA, A)
            pw = {}

            def power(which):
                if which not in pw:
                    h = dc.matrix()
                    if which == "a2":
                        dc.mul(h, A, A)
                    elif which == "a4":
                        dc.mul(h, power("a2"), power("a2"))
                    else:
                        dc.mul(h, power("a2"), power("a4"))
                    pw[which] = h
                return pw[which]

            def nonneg(h):               # expm.cr:205: the reference tests the real parts only
                if cplx:
                    return bool(np.all(dc.download(h, n, n).real >= 0))
                dc.absolute(t1, h)
                dc.axpby(t1, -1.0, h, 1.0)           # |X| - X
                return dc.norm(t1, "M") == 0.0

            def norm_am_of(h, m, scale=1.0, known_nonneg=False):
                src = h
                if not known_nonneg and not nonneg(h):
                    dc.absolute(t2, h)
                    src = t2
                dc.fill(e1, 1.0, 1.0, cols=1)
                x, y = e1, e2
                for _ in range(m):
                    dc.mul(y, src, x, cols=1, trans_a=True)      # e <- X^T e (plain transpose)
                    x, y = y, x
                return dc.norm(x, "M", cols=1) / scale ** m

            norm1_a = dc.norm(A, "O")
            if norm1_a == 0.0:
                return np.asfortranarray(np.eye(n, dtype=dt))
            deg, s = self._expm_plan(lambda w, m: norm_am_of(power(w), m),
                                     lambda m, scale: norm_am_of(absA, m, scale, True), norm1_a)
            c = self._EXPM_PADE[deg]
            sc = [1.0 / 2.0 ** (k * s) for k in range(7)]       # a^k is scaled by 2^-(k s)
            U, V = t1, t2
            if deg == 13:
                a2, a4, a6 = power("a2"), power("a4"), power("a6")
                W = t3
                dc.axpby(W, c[13] * sc[6], a6)                    # c14 a6 + c12 a4 + c10 a2 (scaled)
                dc.axpby(W, c[11] * sc[4], a4, 1.0)
                dc.axpby(W, c[9] * sc[2], a2, 1.0)
                Sa6 = dc.matrix()
                dc.axpby(Sa6, sc[6], a6)
                dc.mul(V, Sa6, W)                                 # V = a6 (...)
                dc.axpby(V, c[7] * sc[6], a6, 1.0)
                dc.axpby(V, c[5] * sc[4], a4, 1.0)
                dc.axpby(V, c[3] * sc[2], a2, 1.0)
                dc.fill(W, 0.0, c[1])
                dc.axpby(V, 1.0, W, 1.0)                          # + c2 I
                dc.axpby(W, sc[1], A)                             # scaled A
                dc.mul(U, W, V)                                   # U = a (...)
                dc.axpby(W, c[12] * sc[6], a6)                    # c13 a6 + c11 a4 + c9 a2
                dc.axpby(W, c[10] * sc[4], a4, 1.0)
                dc.axpby(W, c[8] * sc[2], a2, 1.0)
                dc.mul(V, Sa6, W)
                dc.axpby(V, c[6] * sc[6], a6, 1.0)
                dc.axpby(V, c[4] * sc[4], a4, 1.0)
                dc.axpby(V, c[2] * sc[2], a2, 1.0)
                dc.fill(W, 0.0, c[0])
                dc.axpby(V, 1.0, W, 1.0)
            else:
                m2 = (deg + 3) // 2
                eye = dc.matrix()
                dc.fill(eye, 0.0, 1.0)
                pws = [eye, power("a2")]
                if deg >= 5:
                    pws.append(power("a4"))
                if deg >= 7:
                    pws.append(power("a6"))
                while len(pws) < m2:
                    h = dc.matrix()
                    dc.mul(h, pws[-1], pws[1])
                    pws.append(h)
                W = t3
                dc.fill(W, 0.0, 0.0)
                for j in range(deg + 1, 1, -2):
                    dc.axpby(W, c[j - 1], pws[j // 2 - 1], 1.0)
                dc.mul(U, A, W)
                dc.fill(V, 0.0, 0.0)
                for j in range(deg, 0, -2):
                    dc.axpby(V, c[j - 1], pws[(j + 1) // 2 - 1], 1.0)
            D, N = t3, dc.matrix()
            dc.axpby(D, 1.0, V)
            dc.axpby(D, -1.0, U, 1.0)                             # V - U
            dc.axpby(N, 1.0, V)
            dc.axpby(N, 1.0, U, 1.0)                              # U + V
            dc.solve(D, N)                                        # F = (V - U)^-1 (U + V)
            F, G2 = N, t1
            for _ in range(s):
                dc.mul(G2, F, F)
                F, G2 = G2, F
            return dc.download(F, n, n)

    # matfun.cr:20-56
    def cosm(self):
        if np.iscomplexobj(self.raw):
            return GeneralMatrix(0.5 * (GeneralMatrix(1j * self.raw).expm().raw +
                                        GeneralMatrix(-1j * self.raw).expm().raw))
        return GeneralMatrix(np.asfortranarray(
            GeneralMatrix((1j * self.raw).astype(np.complex128)).expm().raw.real.astype(self.dtype)))

    def sinm(self):
        if np.iscomplexobj(self.raw):
            return GeneralMatrix(-0.5j * (GeneralMatrix(1j * self.raw).expm().raw -
                                          GeneralMatrix(-1j * self.raw).expm().raw))
        return GeneralMatrix(np.asfortranarray(
            GeneralMatrix((1j * self.raw).astype(np.complex128)).expm().raw.imag.astype(self.dtype)))

    def tanm(self):
        return self.sinm() * self.cosm().inv_()

    def coshm(self):
        return GeneralMatrix(0.5 * (self.expm().raw + GeneralMatrix(-self.raw).expm().raw))

    def sinhm(self):
        return GeneralMatrix(0.5 * (self.expm().raw - GeneralMatrix(-self.raw).expm().raw))

    def tanhm(self):
        return self.sinhm() * self.coshm().inv_()

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
            # adjust_symmetric: mirror the computed (upper) triangle, block by block (plain slices:
            # index arrays over the whole triangle cost more than the inversion itself)
            bs = 512
            for j0 in range(0, n, bs):
                j1 = min(n, j0 + bs)
                d = self.raw[j0:j1, j0:j1]
                d[:] = np.triu(d) + np.triu(d, 1).conj().T
                if j1 < n:
                    self.raw[j1:, j0:j1] = self.raw[j0:j1, j1:].conj().T
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


def expm(a): return a.expm()
def cosm(a): return a.cosm()
def sinm(a): return a.sinm()
def tanm(a): return a.tanm()
def coshm(a): return a.coshm()
def sinhm(a): return a.sinhm()
def tanhm(a): return a.tanhm()
def solve(a, b, **kw): return a.solve(b, **kw)
def inv(a): return a.inv()
def solvels(a, b, **kw): return a.solvels(b, **kw)
def lstsq(a, b, method=LSMethod.Auto, **kw): return a.lstsq(b, method, **kw)
def cho_solve(a, b, **kw): return a.cho_solve(b, **kw)


def GMat(data): return GeneralMatrix(data, dtype=np.float64)
def GMat32(data): return GeneralMatrix(data, dtype=np.float32)
def GMatComplex(data): return GeneralMatrix(data, dtype=np.complex128)
