"""The reference's own hot-path specs, restated against the host
mirror `linalg_b200.la`.  Every case is a function of no arguments that uses whatever backend
`la.use_backend()` has installed, so the SAME assertions run on the CPU oracle (tests that are
not marked gpu) and on libla_b200.so (tests marked gpu).

Source of each case: /root/reference/spec/linalg_spec.cr, matrix_spec.cr, power_spec.cr,
flags_spec.cr, benchmarks/multiply.cr (line numbers in the docstrings).
"""
import ctypes as C

import numpy as np

from linalg_b200 import ffi, la
from linalg_b200.la import GMat, GMat32, GMatComplex, GeneralMatrix, LinAlgError, MatrixFlags


def k1_direct_lapack():
    """spec/linalg_spec.cr:5-24 — raw dgetrf_ + dgetri_ incl. lwork=-1 query; exact result."""
    be = la.backend()
    m = np.array([1.0, 0.0, 1.0, 0.0, 4.0, 0.0, 0.0, 0.0, 1.0])
    ipiv = np.zeros(3, dtype=np.int32)
    n, info = C.c_int(3), C.c_int(-99)
    be.lapack("getrf", "d")(C.byref(n), C.byref(n), ffi.ptr(m), C.byref(n), ffi.ptr(ipiv),
                            C.byref(info))
    assert info.value == 0
    assert list(ipiv) == [1, 2, 3]
    size, worksize = C.c_int(-1), np.zeros(1)
    getri = be.lapack("getri", "d")
    getri(C.byref(n), ffi.ptr(m), C.byref(n), ffi.ptr(ipiv), ffi.ptr(worksize), C.byref(size),
          C.byref(info))
    size = C.c_int(int(worksize[0]))
    assert size.value >= 3
    work = np.zeros(size.value)
    getri(C.byref(n), ffi.ptr(m), C.byref(n), ffi.ptr(ipiv), ffi.ptr(work), C.byref(size),
          C.byref(info))
    assert info.value == 0
    assert list(m) == [1.0, 0.0, -1.0, 0.0, 0.25, 0.0, 0.0, 0.0, 1.0]


def k2_inverse_identity():
    """spec/linalg_spec.cr:26-48, :82-89, :108-122 — A * inv(A) eq identity, d / s / z."""
    for ctor, dt in ((GMat, np.float64), (GMat32, np.float32)):
        a = ctor([[1, 0, 1], [0, 4, 0], [0, 0, 1]])
        assert (a * a.inv()) == GeneralMatrix.identity(3, dt)
    a = GMatComplex([[1 + 1j, 0, 1], [0, 4, 0], [0, 0, 1]])
    assert (a * a.inv()) == GeneralMatrix.identity(3, np.complex128)
    # :91-95 submatrix variant
    m4 = np.ones((4, 4)) + np.diag([1.0, 2, 3, 4])
    m3 = GMat(m4[:3, :3])
    assert (m3 * m3.inv()).almost_eq(GeneralMatrix.identity(3))


def k3_singular_raises():
    """spec/linalg_spec.cr:97-106 — singular matrix -> LinAlgError (getrf info = 2)."""
    a = GMat([[1, 0, 1], [0, 0, 0], [0, 0, 1]])
    try:
        a.inv_()
    except LinAlgError as e:
        assert "returned 2" in str(e)
    else:
        raise AssertionError("LinAlgError not raised")


def k4_posdef_inverse():
    """spec/linalg_spec.cr:124-136 — detect => PositiveDefinite via potrf; potrf+potri exact."""
    a = GMat([[1, 1, 1], [1, 2, 3], [1, 3, 6]])
    a.detect()
    assert a.flags.positive_definite and a.flags.symmetric
    assert a.inv() == GMat([[3.0, -3.0, 1.0], [-3.0, 5.0, -2.0], [1.0, -2.0, 1.0]])


def k5_solve_f32():
    """spec/linalg_spec.cr:138-144, :270-277 — sgesv and sgetrf+sgetrs eq inv*b."""
    a = GMat32([[2, 4], [2, 8]])
    b = GMat32([[2], [4]])
    assert la.solve(a, b) == (a.inv() * b)
    assert a.lu_factor().solve(b) == (a.inv() * b)


def k6_posv():
    """spec/linalg_spec.cr:146-168 — posv path, real and complex, |A x - b| < 1e-6."""
    for ctor in (GMat, GMatComplex):
        a = ctor([[1, 0, 0, 0], [4, 5, 1, 0], [0, 1, -1, 0], [0, 0, 2, 1]])
        a = a * a.t()
        assert a.detect_(MatrixFlags.PositiveDefinite)
        b = ctor([[1, 2, 7, 1]]).t()
        x = a.solve(b)
        assert np.all(np.abs((a * x).raw - b.raw) < 1e-6)


def k7_det():
    """spec/linalg_spec.cr:170-182 — det via getrf eq 9; complex triangular eq 5."""
    a = GMat([[1, 2, 3], [4, 5, 7], [-1, 1, -1]])
    assert a.det() == 9
    c = GMatComplex([[1j, 2, 3], [0, 5, 7], [0, 0, -1j]])
    assert c.det() == 5


def k8_least_squares_wide():
    """spec/linalg_spec.cr:183-231 — solvels / lstsq(QR) on the wide 2x3 system (gelqf branch)."""
    x_oct = np.array([[0.918032], [3.54098], [1.27869]])
    a, b = GMat32([[1, 2, 0], [0, 4, 3]]), GMat32([[8], [18]])
    assert np.all(np.abs(la.solvels(a, b).raw - x_oct) < 1e-3)
    for ctor in (GMat, GMatComplex):
        a, b = ctor([[1, 2, 0], [0, 4, 3]]), ctor([[8], [18]])
        x, rank, s = la.lstsq(a, b, la.LSMethod.QR)
        assert np.all(np.abs(x.raw - x_oct) < 1e-5)
        assert s == []


def k9_lu_rectangular():
    """spec/linalg_spec.cr:256-268 — lu of 5x4 Float32: p*l*u ~ a, L lower, U upper."""
    a = GMat32([[1, 2, 3, 4], [5, 6, 7, 8], [1, 2, 10, 0], [50, 6, 7, 8], [1, 21, 1, 0]])
    p, l, u = a.lu()
    assert (p * l * u).almost_eq(a)
    assert l.detect_(MatrixFlags.LowerTriangular)
    assert not l.detect_(MatrixFlags.UpperTriangular)
    assert u.detect_(MatrixFlags.UpperTriangular)


def k10_cholesky_complex():
    """spec/linalg_spec.cr:279-293 — complex cholesky(lower), cho_solve from L and U equal."""
    a = GMatComplex([[1, -2j], [2j, 5]])
    c = a.cholesky(lower=True)
    assert (c * c.conjt()).almost_eq(a)
    chol1 = a.cholesky(lower=True, dont_clean=True)
    chol2 = a.cholesky(lower=False, dont_clean=False)
    b = GMatComplex([[2], [4]])
    x1, x2 = chol1.cho_solve(b), chol2.cho_solve(b)
    assert x1 == x2
    assert (a * x1 - b).almost_eq(GeneralMatrix.zeros(2, 1, np.complex128))


def k11_qr_complex():
    """spec/linalg_spec.cr:353-365 — complex qr: Q unitary, q*r ~ a, qr_r ~ r."""
    a = GMatComplex([[12 + 1j, -51, 4], [6, 167 - 2j, -68], [-4, 24, -41 + 3j]])
    q, r = a.qr()
    assert q.detect_(MatrixFlags.Orthogonal)
    assert (q * r).almost_eq(a)
    assert a.qr_r().almost_eq(r)
    assert r.detect_(MatrixFlags.UpperTriangular)


def k12_k13_products():
    """spec/matrix_spec.cr:58-64, :523-526 — small exact products; size mismatch raises."""
    m1 = GMat(np.fromfunction(lambda i, j: i + j, (3, 4)))   # GMat.new(3, 4) { |i, j| i + j }
    m2 = GMat(np.fromfunction(lambda i, j: i - j, (4, 2)))   # GMat.new(4, 2) { |i, j| i - j }
    assert (m1 * m2) == GMat([[14, 8], [20, 10], [26, 12]])  # row-major listing in the spec
    try:
        m1 * GMat([[1, 2], [3, 4]])
    except ValueError:
        pass
    else:
        raise AssertionError("size mismatch not raised")
    for dt in (np.float64, np.float32, np.complex128):
        o = GeneralMatrix.ones(3, 4, dt) * GeneralMatrix.ones(4, 3, dt)
        assert o == GeneralMatrix.wrap(np.asfortranarray(4 * np.ones((3, 3), dtype=dt)))


def k14_power():
    """spec/power_spec.cr:16-19 — m**5 eq m*m*m*m*m; m**(-2) eq (m*m).inv (small ints)."""
    m = GMat([[1, 2], [3, 4]])
    assert (m ** 5) == (m * m * m * m * m)
    assert (m ** (-2)).almost_eq((m * m).inv())


def k15_trmm_flags():
    """spec/flags_spec.cr:97-101 — triangular x triangular goes through trmm, keeps the flag."""
    a = GMat([[1, 2, 3], [0, 4, 5], [0, 0, 6]]).assume_(MatrixFlags.UpperTriangular)
    b = GMat([[2, 1, 0], [0, 1, 1], [0, 0, 3]]).assume_(MatrixFlags.UpperTriangular)
    c = a * b
    assert c.flags.upper_triangular
    assert c == GMat(np.array([[1, 2, 3], [0, 4, 5], [0, 0, 6.0]]) @
                     np.array([[2, 1, 0], [0, 1, 1], [0, 0, 3.0]]))
    lo = a.t()
    d = lo * GMat([[1, 2, 3], [4, 5, 6], [7, 8, 9]])
    assert d == GMat(lo.raw @ np.array([[1, 2, 3], [4, 5, 6], [7, 8, 9.0]]))


def k17_multiply_selfchecks():
    """benchmarks/multiply.cr:77-85 — a*a.inv, ge/sy/tr products ~ I on 6x6."""
    rng = np.random.default_rng(17)
    g = rng.random((6, 6)) + 6 * np.eye(6)
    a = GMat(g)
    eye = GeneralMatrix.identity(6)
    assert (a * a.inv()).almost_eq(eye)
    s = GMat(g + g.T).assume_(MatrixFlags.Symmetric)
    assert (s * s.inv()).almost_eq(eye, 1e-9)        # symm path
    t = GMat(np.triu(g)).assume_(MatrixFlags.UpperTriangular)
    assert (t * t.inv()).almost_eq(eye, 1e-9)        # trmm + trtri path
    h = rng.random((6, 6)) + 1j * rng.random((6, 6))
    hm = GMatComplex(h + h.conj().T + 12 * np.eye(6)).assume_(MatrixFlags.Hermitian)
    assert (hm * hm.inv()).almost_eq(GeneralMatrix.identity(6, np.complex128), 1e-9)  # hemm


ALL_CASES = [k1_direct_lapack, k2_inverse_identity, k3_singular_raises, k4_posdef_inverse,
             k5_solve_f32, k6_posv, k7_det, k8_least_squares_wide, k9_lu_rectangular,
             k10_cholesky_complex, k11_qr_complex, k12_k13_products, k14_power, k15_trmm_flags,
             k17_multiply_selfchecks]
