"""GPU parity tests of the SURVEY.md §8(f) rows built after the hot path: RQ / LQ / QL
(src/linalg/rq_lq_ql.cr), pivoted QR (src/linalg/qr.cr:38-41), the matrix norms
(src/linalg/linalg.cr:721-748), ?lacpy / ?laset, and the device-matrix handles behind `**`
(src/linalg/power.cr:19-30).  Everything goes through the C ABI with host pointers and is
compared with the CPU oracle (the library the reference links) on the same inputs."""
import ctypes as C

import numpy as np
import pytest

from linalg_b200 import ffi, la
from oracle import checks

pytestmark = pytest.mark.gpu

DTYPES = [np.float32, np.float64, np.complex128]
IDS = ["s", "d", "z"]


def _rand(rng, shape, dt, lo=-0.5):
    a = rng.random(shape) + lo
    if np.dtype(dt) == np.complex128:
        a = a + 1j * (rng.random(shape) + lo)
    return np.asfortranarray(a.astype(dt))


def _tol(dt):
    return 2e-4 if np.dtype(dt) == np.float32 else 1e-11


def _both(fn, product_abi, oracle_abi):
    with la.use_backend(product_abi):
        got = fn()
    with la.use_backend(oracle_abi):
        ref = fn()
    return got, ref


@pytest.mark.parametrize("dt", DTYPES, ids=IDS)
@pytest.mark.parametrize("fam", ["gerqf", "gelqf", "geqlf"])
@pytest.mark.parametrize("shape", [(40, 40), (70, 30), (30, 70), (300, 129), (1, 5), (5, 1)])
def test_rq_lq_ql_factors_have_lapack_storage(fam, dt, shape, product_abi, oracle_abi):
    """Same A (factor + reflectors) and tau as LAPACK, element by element."""
    rng = np.random.default_rng(21)
    A = _rand(rng, shape, dt)
    m, n = shape

    def run():
        a = A.copy(order="F")
        tau = np.zeros(min(m, n), dtype=dt)
        la._lapack(fam, ffi.LETTER[np.dtype(dt)], m, n, a, m, tau)
        return a, tau
    (a1, t1), (a0, t0) = _both(run, product_abi, oracle_abi)
    assert checks.rel(a1 - a0, a0) < _tol(dt)
    assert np.abs(t1 - t0).max() < _tol(dt) * 10


@pytest.mark.parametrize("dt", DTYPES, ids=IDS)
@pytest.mark.parametrize("letters,shape", [("rq", (40, 40)), ("rq", (129, 129)), ("lq", (64, 64)),
                                           ("lq", (5, 5)), ("ql", (33, 33)), ("ql", (200, 200))])
def test_rq_lq_ql_decompositions_multiply_back(letters, dt, shape, product_abi, oracle_abi):
    """The reference's rq / lq / ql (factor + org??) on square matrices (what its `triu!` / `tril!`
    of the factor is meaningful for): orthonormal Q, triangular factor, product = A, and the
    same factors as with the oracle."""
    rng = np.random.default_rng(22)
    A = la.GeneralMatrix(_rand(rng, shape, dt))

    def run():
        x, y = getattr(A, letters)()
        return x.raw.copy(), y.raw.copy()
    (x1, y1), (x0, y0) = _both(run, product_abi, oracle_abi)
    assert checks.rel(x1 @ y1 - A.raw, A.raw) < _tol(dt)
    assert checks.rel(x1 - x0, x0) < _tol(dt) * 10 and checks.rel(y1 - y0, y0) < _tol(dt) * 10
    q = x1 if letters[0] == "q" else y1
    k = min(shape)
    g = q.conj().T @ q if letters == "ql" else q @ q.conj().T
    assert np.abs(g - np.eye(k)).max() < _tol(dt) * 10


@pytest.mark.parametrize("dt", DTYPES, ids=IDS)
@pytest.mark.parametrize("shape", [(60, 40), (40, 60), (200, 200), (500, 37), (1, 1)])
def test_pivoted_qr_has_lapack_pivots_and_factors(dt, shape, product_abi, oracle_abi):
    rng = np.random.default_rng(23)
    m, n = shape
    # column norms well separated, so that the greedy pivot choice is unambiguous
    A = la.GeneralMatrix(np.asfortranarray(_rand(rng, shape, dt) * (1.0 + np.arange(n) % 7).astype(dt)))

    def run():
        a, tau, pvt = A.qr_raw(pivoting=True)
        return a.raw.copy(), tau.copy(), pvt.copy()
    (a1, t1, p1), (a0, t0, p0) = _both(run, product_abi, oracle_abi)
    assert np.array_equal(p1, p0)
    assert sorted(p1) == list(range(1, n + 1))
    assert checks.rel(a1 - a0, a0) < _tol(dt) * 10
    assert np.abs(t1[:min(m, n)] - t0[:min(m, n)]).max() < _tol(dt) * 10
    d = np.abs(np.diagonal(a1))
    assert np.all(d[:-1] >= d[1:] * (1 - 1e-6))                   # |r_11| >= |r_22| >= ...


def test_pivoted_qr_keeps_fixed_columns_in_front_and_finds_the_rank(product_abi, oracle_abi):
    rng = np.random.default_rng(24)
    m, n, r = 80, 50, 17
    A = np.asfortranarray(rng.random((m, r)) @ rng.random((r, n)))
    with la.use_backend(product_abi):
        assert la.GeneralMatrix(A).rank(1e-9 * np.abs(A).max() * n, method=la.RankMethod.QRP) == r
    L = "d"

    def run(abi):
        a = A.copy(order="F")
        jpvt = np.zeros(n, dtype=np.int32)
        jpvt[[4, 9, 30]] = 1                                       # fixed columns
        tau = np.zeros(min(m, n))
        with la.use_backend(abi):
            la._lapack("geqp3", L, m, n, a, m, jpvt, tau)
        return a, jpvt, tau
    a1, p1, t1 = run(product_abi)
    a0, p0, t0 = run(oracle_abi)
    assert sorted(p1[:3]) == [5, 10, 31]
    assert np.array_equal(p1[:r], p0[:r])                          # beyond the rank: rounding noise
    assert np.abs(np.triu(a1)[:r, :r] - np.triu(a0)[:r, :r]).max() < 1e-10


@pytest.mark.parametrize("dt", DTYPES, ids=IDS)
@pytest.mark.parametrize("kind", list(la.MatrixNorm))
def test_norms_match_lapack_for_every_shape_flag(dt, kind, product_abi, oracle_abi):
    rng = np.random.default_rng(25)
    n = 301
    G = _rand(rng, (n, n), dt)
    H = np.asfortranarray(G + G.conj().T)
    S = np.asfortranarray(G + G.T)
    cases = [la.GeneralMatrix(_rand(rng, (257, 130), dt)),
             la.GeneralMatrix(S, flags=la.MatrixFlags.Symmetric),
             la.GeneralMatrix(np.triu(G), flags=la.MatrixFlags.UpperTriangular),
             la.GeneralMatrix(np.tril(G), flags=la.MatrixFlags.LowerTriangular)]
    if np.dtype(dt) == np.complex128:
        cases.append(la.GeneralMatrix(H, flags=la.MatrixFlags.Hermitian | la.MatrixFlags.Symmetric))
    for mtx in cases:
        got, ref = _both(lambda: mtx.norm(kind), product_abi, oracle_abi)
        assert abs(got - ref) <= (1e-5 if np.dtype(dt) == np.float32 else 1e-13) * abs(ref)


@pytest.mark.parametrize("dt", DTYPES, ids=IDS)
@pytest.mark.parametrize("uplo", [b"U", b"L", b"A"])
def test_lacpy_and_laset_match_lapack(dt, uplo, product_abi, oracle_abi):
    rng = np.random.default_rng(26)
    L = ffi.LETTER[np.dtype(dt)]
    for m, n in [(7, 5), (5, 7), (1500, 1100)]:                   # the last one uses host threads
        A = _rand(rng, (m + 3, n), dt)
        B0 = _rand(rng, (m + 1, n), dt)
        out = []
        for abi in (product_abi, oracle_abi):
            B = B0.copy(order="F")
            abi.lapack("lacpy", L)(uplo, la._ci(m), la._ci(n), ffi.ptr(A), la._ci(m + 3), ffi.ptr(B),
                                   la._ci(m + 1))
            Z = B0.copy(order="F")
            al = np.array([2.5], dtype=dt)
            be = np.array([-1.25], dtype=dt)
            abi.lapack("laset", L)(uplo, la._ci(m), la._ci(n), ffi.ptr(al), ffi.ptr(be), ffi.ptr(Z),
                                   la._ci(m + 1))
            out.append((B, Z))
        assert np.array_equal(out[0][0], out[1][0]) and np.array_equal(out[0][1], out[1][1])


@pytest.mark.parametrize("dt", DTYPES, ids=IDS)
def test_power_on_resident_operands_is_bit_identical_to_the_pointer_path(dt, product_abi):
    """K14 (`**`) through the handles: one upload, products on resident matrices, one download."""
    rng = np.random.default_rng(27)
    n = 160
    A = la.GeneralMatrix(_rand(rng, (n, n), dt) / np.sqrt(n))
    with la.use_backend(product_abi):
        for p in (2, 5, 8, 13):
            fast = A ** p
            # the pointer path: the same repeated squaring, every product through cblas_?gemm
            result, base, first, q = None, A, True, p
            while q > 0:
                if q & 1:
                    result = base.clone() if first else result * base
                    first = False
                q >>= 1
                if q:
                    base = base * base
            assert np.array_equal(fast.raw, result.raw), p
        ref = np.linalg.matrix_power(A.raw.astype(np.complex128 if np.iscomplexobj(A.raw) else np.float64), 5)
        assert checks.rel((A ** 5).raw - ref, ref) < _tol(dt)


def test_handles_report_their_geometry_and_refuse_bad_arguments(product_abi):
    cd = product_abi.cdll
    cd.la_b200_mat_create.restype = C.c_void_p
    cd.la_b200_mat_create.argtypes = [C.c_int, C.c_int64, C.c_int64]
    for f in (cd.la_b200_mat_ld, cd.la_b200_mat_rows, cd.la_b200_mat_cols):
        f.restype, f.argtypes = C.c_int64, [C.c_void_p]
    cd.la_b200_mat_free.argtypes = [C.c_void_p]
    cd.la_b200_mat_upload.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
    cd.la_b200_mat_download.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
    assert cd.la_b200_mat_create(ord("q"), 4, 4) is None
    h = C.c_void_p(cd.la_b200_mat_create(ord("d"), 37, 5))
    assert (cd.la_b200_mat_rows(h), cd.la_b200_mat_cols(h), cd.la_b200_mat_ld(h)) == (37, 5, 48)
    a = np.asfortranarray(np.arange(40 * 5, dtype=np.float64).reshape(40, 5, order="F"))
    b = np.zeros((39, 5), order="F")
    assert cd.la_b200_mat_upload(h, ffi.ptr(a), 36) == -3
    assert cd.la_b200_mat_upload(h, ffi.ptr(a), 40) == 0
    assert cd.la_b200_mat_download(h, ffi.ptr(b), 39) == 0
    assert np.array_equal(b[:37], a[:37]) and not b[37:].any()
    cd.la_b200_mat_free(h)


# ---- matrix functions on top of the path (src/linalg/expm.cr, matfun.cr) -------
from oracle import expm_ref  # noqa: E402


def test_expm_reference_spec_vectors(product_abi):
    """spec/expm_spec.cr and spec/matfun_spec.cr through the product (special cases, the 2 x 2
    closed forms and the small general path)."""
    with la.use_backend(product_abi):
        assert np.array_equal(la.GMat(np.zeros((2, 2))).expm().raw, np.eye(2))
        assert np.array_equal(la.GMat(np.zeros((80, 80))).expm().raw, np.eye(80))     # resident chain
        assert abs(la.GMat(np.ones((1, 1))).expm().raw[0, 0] - np.e) < 1e-14
        m = la.GMat([[0, 6, 0, 0], [0, 0, 6, 0], [0, 0, 0, 6], [0, 0, 0, 0]])
        mres = np.array([[1.0, 6.0, 18.0, 36.0], [0.0, 1.0, 6.0, 18.0], [0.0, 0.0, 1.0, 6.0], [0.0, 0.0, 0.0, 1.0]])
        assert np.abs(m.expm().raw - mres).max() < 1e-12
        assert np.abs(la.GMat32(m.raw).expm().raw - mres).max() < 1e-4
        sym = la.GMat(m.raw + m.raw.T).expm().raw
        assert abs(sym[0, 0] - 2288.37752914804) < 1e-8 and abs(sym[1, 2] - 5946.76670170141) < 1e-8
        bad = la.GMat([[1, 1e9], [0, 1]]).expm().raw
        assert np.allclose(bad, [[np.e, np.e * 1e9], [0, np.e]], rtol=1e-14)
        z = la.GMatComplex([[0, 1j, 1], [2, 0, -1j], [0, 0, 1]]).expm().raw
        assert abs(z[0, 0] - (0.83373 + 0.98890j)) < 1e-4 and abs(z[1, 2] - (1.85841 - 1.47252j)) < 1e-4
        d = la.GeneralMatrix(np.diag([np.pi * 1j, 5, 1]).astype(np.complex128), flags=la.MatrixFlags.Triangular)
        assert np.allclose(d.expm().raw, np.diag([-1, np.e ** 5, np.e]), rtol=1e-14, atol=1e-14)
        c = la.GMatComplex([[1.0, 2.0], [-1.0, 3.0]])
        assert np.abs(la.expm(la.GMatComplex(1j * c.raw)).raw -
                      np.array([[0.42645930 + 1.89217551j, -2.13721484 - 0.97811252j],
                                [1.06860742 + 0.48905626j, -1.71075555 + 0.91406299j]])).max() < 1e-5
        assert np.allclose(la.cosm(c).raw + 1j * la.sinm(c).raw, la.expm(la.GMatComplex(1j * c.raw)).raw, atol=1e-13)
        a = la.GMat([[1.0, 3.0], [1.0, 4.0]])
        assert np.abs(la.sinhm(a).raw - np.array([[10.57300653, 39.28826594], [13.09608865, 49.86127247]])).max() < 1e-6
        assert np.allclose(la.tanhm(a).raw, la.sinhm(a).raw @ np.linalg.inv(la.coshm(a).raw), rtol=1e-11)


@pytest.mark.parametrize("dt", DTYPES, ids=IDS)
@pytest.mark.parametrize("n,scale", [(5, 0.05), (9, 0.6), (24, 2.0), (40, 12.0),      # host chain
                                     (64, 0.02), (100, 1.5), (200, 30.0), (333, 8.0)])  # resident chain
def test_expm_matches_the_oracle_over_all_pade_degrees(dt, n, scale, product_abi):
    """The product's expm (host chain below order 64, device-resident chain from there) against
    the CPU restatement of the reference's algorithm on the same input."""
    rng = np.random.default_rng(31 + n)
    a = _rand(rng, (n, n), dt) * (scale / np.sqrt(n))
    ref = expm_ref.expm(a.astype(np.complex128 if np.iscomplexobj(a) else np.float64))
    with la.use_backend(product_abi):
        got = la.GeneralMatrix(a).expm().raw
    tol = 2e-4 if np.dtype(dt) == np.float32 else 1e-11
    assert np.linalg.norm(got - ref) / np.linalg.norm(ref) < tol


def test_expm_resident_chain_equals_the_host_chain(product_abi):
    rng = np.random.default_rng(77)
    n = 150
    for dt, scale in ((np.float64, 20.0), (np.complex128, 3.0)):
        a = la.GeneralMatrix(_rand(rng, (n, n), dt) * (scale / np.sqrt(n)))
        with la.use_backend(product_abi):
            host = a._expm_host()
            res = a._expm_resident()
        assert np.linalg.norm(host - res) / np.linalg.norm(host) < 1e-12


@pytest.mark.parametrize("dt", DTYPES, ids=IDS)
def test_geadd_and_geabs_on_resident_matrices(dt, product_abi):
    rng = np.random.default_rng(41)
    n = 77
    a, b = _rand(rng, (n, n), dt), _rand(rng, (n, n), dt)
    with la.use_backend(product_abi):
        with la._DeviceChain(ffi.LETTER[np.dtype(dt)], n, dt) as dc:
            ha, hb = dc.matrix(), dc.matrix()
            dc.upload(ha, a)
            dc.upload(hb, b)
            dc.axpby(hb, 2.5, ha, -0.75)
            got = dc.download(hb, n, n)
            assert np.allclose(got, 2.5 * a - 0.75 * b, rtol=0, atol=2e-6 if np.dtype(dt) == np.float32 else 1e-15)
            dc.axpby(hb, -1.0, ha)                     # beta = 0: B is not read (NaN-safe)
            assert np.array_equal(dc.download(hb, n, n), -a)
            dc.absolute(hb, ha)
            absref = np.abs(a).astype(dt)
            assert np.allclose(dc.download(hb, n, n), absref, rtol=1e-6 if np.dtype(dt) == np.float32 else 1e-15)
            assert abs(dc.norm(ha, "O") - np.abs(a).sum(axis=0).max()) < 1e-4 * n
