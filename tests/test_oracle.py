"""CPU tests (not gpu): pin the oracles.

1. The reference's golden vectors (K1-K17) hold on the OpenBLAS/LAPACK oracle through the host
   mirror — this validates the mirror and the ABI binding themselves.
2. The plain-C restatement (oracle/la_ref.c) agrees with OpenBLAS/LAPACK on seeded inputs,
   including identical pivot vectors, Householder conventions (R, tau) and info codes.
"""
import numpy as np
import pytest

from linalg_b200 import la
from oracle import c_ref, checks
from tests import spec_cases

DTYPES = [np.float32, np.float64, np.complex128]


def _rand(rng, shape, dt):
    a = rng.random(shape)
    if np.dtype(dt) == np.complex128:
        a = a + 1j * rng.random(shape)
    return np.asfortranarray(a.astype(dt))


@pytest.mark.parametrize("case", spec_cases.ALL_CASES, ids=lambda f: f.__name__)
def test_reference_specs_on_openblas(case, oracle_abi):
    with la.use_backend(oracle_abi):
        case()


@pytest.mark.parametrize("dt", DTYPES)
def test_c_ref_gemm_matches_numpy(rng, dt):
    A, B, Cm = _rand(rng, (7, 5), dt), _rand(rng, (5, 9), dt), _rand(rng, (7, 9), dt)
    got = c_ref.gemm("N", "N", 1.5, A, B, -0.5, Cm)
    ref = 1.5 * (A.astype(np.complex128) @ B) - 0.5 * Cm
    assert checks.rel(got - ref, ref) < 50 * np.finfo(np.dtype(dt)).eps
    got = c_ref.gemm("C", "T", 1.0, _rand(rng, (5, 7), dt), _rand(rng, (9, 5), dt), 0.0, Cm)
    assert got.shape == (7, 9)


@pytest.mark.parametrize("dt", DTYPES)
def test_c_ref_lu_matches_lapack(rng, dt, oracle_abi):
    for shape in [(40, 40), (50, 30), (30, 50)]:
        A = _rand(rng, shape, dt)
        LU, ipiv, info = c_ref.getrf(A)
        assert info == 0
        with la.use_backend(oracle_abi):
            ref = la.GeneralMatrix(A)
            rp = np.zeros(min(shape), dtype=np.int32)
            la._lapack("getrf", ref.letter, shape[0], shape[1], ref.raw, shape[0], rp)
        assert list(ipiv) == list(rp)
        assert checks.rel(LU - ref.raw, ref.raw) < 1e3 * np.finfo(np.dtype(dt)).eps
        assert checks.lu_residual(A, LU, ipiv) < checks.tol(dt)


@pytest.mark.parametrize("dt", DTYPES)
@pytest.mark.parametrize("uplo", ["L", "U"])
def test_c_ref_cholesky_matches_lapack(rng, dt, uplo, oracle_abi):
    n = 33
    G = _rand(rng, (n, n), dt)
    A = np.asfortranarray((G @ G.conj().T + n * np.eye(n)).astype(dt))
    F, info = c_ref.potrf(uplo, A)
    assert info == 0
    with la.use_backend(oracle_abi):
        ref = la.GeneralMatrix(A).cholesky_(lower=(uplo == "L"), dont_clean=True)
    tri = np.tril if uplo == "L" else np.triu
    assert checks.rel(tri(F) - tri(ref.raw), tri(ref.raw)) < 1e3 * np.finfo(np.dtype(dt)).eps
    # the other triangle is untouched
    other = np.triu(A, 1) if uplo == "L" else np.tril(A, -1)
    assert np.array_equal(np.triu(F, 1) if uplo == "L" else np.tril(F, -1), other)
    # not positive definite -> info = order of the failing minor
    A2 = A.copy()
    A2[5, 5] = -1.0
    _, info = c_ref.potrf(uplo, A2)
    assert info == 6


@pytest.mark.parametrize("dt", DTYPES)
def test_c_ref_qr_conventions_match_lapack(rng, dt, oracle_abi):
    for shape in [(30, 12), (12, 12), (10, 20)]:
        A = _rand(rng, shape, dt) - 0.5
        QR, tau = c_ref.geqrf(A)
        with la.use_backend(oracle_abi):
            ref, rtau = la.GeneralMatrix(A).qr_raw()
        scale = 1e3 * np.finfo(np.dtype(dt)).eps
        assert checks.rel(QR - ref.raw, ref.raw) < scale
        assert checks.rel(tau - rtau[: len(tau)], rtau[: len(tau)]) < scale


@pytest.mark.parametrize("dt", DTYPES)
@pytest.mark.parametrize("shape", [(20, 8), (8, 20), (9, 9)])
def test_c_ref_gels_matches_lapack(rng, dt, shape, oracle_abi):
    A = _rand(rng, shape, dt)
    B = _rand(rng, (shape[0], 3), dt)
    _, X, info = c_ref.gels("N", A, B)
    assert info == 0
    with la.use_backend(oracle_abi):
        ref = la.solvels(la.GeneralMatrix(A), la.GeneralMatrix(B))
    n = shape[1]
    assert checks.rel(X[:n] - ref.raw[:n], ref.raw[:n]) < 1e4 * np.finfo(np.dtype(dt)).eps


def test_pivot_tie_break_is_first_maximum():
    # two rows with the same |a|: i?amax takes the first one
    A = np.asfortranarray(np.array([[1.0, 2.0], [-1.0, 5.0], [1.0, 7.0]]))
    _, ipiv, _ = c_ref.getrf(A)
    assert ipiv[0] == 1
    Z = np.asfortranarray(np.array([[3 + 4j, 1], [5 + 2j, 2], [0, 1]], dtype=np.complex128))
    _, ipiv, _ = c_ref.getrf(Z)  # cabs1: |3|+|4| = 7 = |5|+|2|
    assert ipiv[0] == 1
