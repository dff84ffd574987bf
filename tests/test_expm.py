"""The matrix-function oracle (oracle/expm_ref.py) against the reference's own golden vectors
(spec/expm_spec.cr, spec/matfun_spec.cr) and against scipy's expm — CPU only."""
import numpy as np
import pytest

from oracle import expm_ref as R

M4 = np.array([[0, 6, 0, 0], [0, 0, 6, 0], [0, 0, 0, 6], [0, 0, 0, 0]], dtype=np.float64)
M4_EXP = np.array([[1.0, 6.0, 18.0, 36.0], [0.0, 1.0, 6.0, 18.0], [0.0, 0.0, 1.0, 6.0], [0.0, 0.0, 0.0, 1.0]])
SYM_EXP = np.array([
    [2288.37752914804, 3687.89823361033, 3669.66066846686, 2258.86846809108],
    [3687.89823361033, 5958.03819761490, 5946.76670170141, 3669.66066846686],
    [3669.66066846686, 5946.76670170141, 5958.03819761491, 3687.89823361034],
    [2258.86846809108, 3669.66066846686, 3687.89823361034, 2288.37752914805]])


def _close(a, b, tol):
    return np.abs(np.asarray(a) - np.asarray(b)).max() <= tol * max(1.0, np.abs(b).max())


def test_expm_spec_vectors_double():           # spec/expm_spec.cr:18-47
    assert np.array_equal(R.expm(np.zeros((2, 2))), np.eye(2))
    assert _close(R.expm(np.ones((1, 1))), [[np.e]], 1e-14)
    assert _close(R.expm(M4), M4_EXP, 1e-13)
    assert _close(R.expm(-M4), M4_EXP * np.array([[1, -1, 1, -1], [1, 1, -1, 1], [1, 1, 1, -1], [1, 1, 1, 1]]), 1e-13)
    assert _close(R.expm(M4 + M4.T), SYM_EXP, 1e-12)
    assert _close(R.expm(np.array([[1, 1e9], [0, 1]])), [[np.e, np.e * 1e9], [0, np.e]], 1e-14)


def test_expm_spec_vectors_single():           # spec/expm_spec.cr:5-16
    assert np.array_equal(R.expm(np.zeros((2, 2), dtype=np.float32)), np.eye(2, dtype=np.float32))
    out = R.expm(M4.astype(np.float32))
    assert out.dtype == np.float32 and _close(out, M4_EXP, 1e-5)


def test_expm_spec_vectors_complex():          # spec/expm_spec.cr:71-81
    m = np.array([[0, 1j, 1], [2, 0, -1j], [0, 0, 1]], dtype=np.complex128)
    ref = np.array([[0.83373 + 0.98890j, -0.33175 + 0.96671j, 2.39036 + 0.53776j],
                    [1.93342 + 0.66349j, 0.83373 + 0.98890j, 1.85841 - 1.47252j],
                    [0, 0, 2.71828]])
    assert np.abs(R.expm(m) - ref).max() < 1e-4
    d = np.diag([np.pi * 1j, 5, 1]).astype(np.complex128)
    assert _close(R.expm(d, diagonal=True), np.diag([-1, np.e ** 5, np.e]), 1e-14)
    assert _close(R.expm(d), np.diag([-1, np.e ** 5, np.e]), 1e-12)      # the general path agrees


def test_matfun_spec_vectors():                # spec/matfun_spec.cr
    m = np.array([[1.0, 2.0], [-1.0, 3.0]], dtype=np.complex128)
    ref = np.array([[0.42645930 + 1.89217551j, -2.13721484 - 0.97811252j],
                    [1.06860742 + 0.48905626j, -1.71075555 + 0.91406299j]])
    assert np.abs(R.expm(1j * m) - ref).max() < 1e-5
    assert _close(R.cosm(m) + 1j * R.sinm(m), R.expm(1j * m), 1e-13)
    assert _close(R.tanm(m), R.sinm(m) @ np.linalg.inv(R.cosm(m)), 1e-13)
    a = np.array([[1.0, 3.0], [1.0, 4.0]])
    assert np.abs(R.sinhm(a) - np.array([[10.57300653, 39.28826594], [13.09608865, 49.86127247]])).max() < 1e-6
    assert _close(R.tanhm(a), R.sinhm(a) @ np.linalg.inv(R.coshm(a)), 1e-13)


@pytest.mark.parametrize("n,scale", [(3, 0.01), (5, 0.2), (8, 1.0), (12, 3.0), (20, 10.0), (33, 40.0)])
@pytest.mark.parametrize("kind", ["real", "complex"])
def test_expm_matches_scipy_over_all_pade_degrees(n, scale, kind):
    sl = pytest.importorskip("scipy.linalg")
    rng = np.random.default_rng(100 + n)
    a = (rng.random((n, n)) - 0.5) * scale / np.sqrt(n)
    if kind == "complex":
        a = a + 1j * (rng.random((n, n)) - 0.5) * scale / np.sqrt(n)
    got, ref = R.expm(a), sl.expm(a)
    assert np.linalg.norm(got - ref) / np.linalg.norm(ref) < 1e-12


# ---- the host mirror (linalg_b200/la.py) through the ORACLE library: the degree selection, the
# norm estimates and the Pade sums of la.py are host logic and are checked here without a GPU
from linalg_b200 import la  # noqa: E402


@pytest.fixture(scope="module")
def cpu_abi():
    from oracle import openblas_ref
    return openblas_ref.abi()


@pytest.mark.parametrize("n,scale", [(3, 0.01), (5, 0.2), (8, 1.0), (12, 3.0), (20, 10.0), (70, 25.0)])
@pytest.mark.parametrize("kind", ["real", "complex", "single"])
def test_la_expm_host_chain_follows_the_reference_algorithm(n, scale, kind, cpu_abi):
    rng = np.random.default_rng(200 + n)
    a = (rng.random((n, n)) - 0.5) * scale / np.sqrt(n)
    if kind == "complex":
        a = a + 1j * (rng.random((n, n)) - 0.5) * scale / np.sqrt(n)
    if kind == "single":
        a = a.astype(np.float32)
    ref = R.expm(a.astype(np.complex128 if np.iscomplexobj(a) else np.float64))
    with la.use_backend(cpu_abi):
        got = la.GeneralMatrix(a).expm().raw
    tol = 5e-4 if kind == "single" else 1e-11
    assert got.dtype == a.dtype
    assert np.linalg.norm(got - ref) / np.linalg.norm(ref) < tol


def test_la_expm_special_cases_and_matfun(cpu_abi):
    with la.use_backend(cpu_abi):
        assert np.array_equal(la.GMat(np.zeros((5, 5))).expm().raw, np.eye(5))
        assert _close(la.GMat(M4).expm().raw, M4_EXP, 1e-13)
        assert _close(la.GMat(M4 + M4.T).expm().raw, SYM_EXP, 1e-12)
        assert _close(la.GMat([[1, 1e9], [0, 1]]).expm().raw, [[np.e, np.e * 1e9], [0, np.e]], 1e-14)
        up = la.GeneralMatrix(np.array([[1.0, 2.0], [0.0, 3.0]]), flags=la.MatrixFlags.UpperTriangular)
        assert _close(up.expm().raw, R.expm(up.raw), 1e-14)                  # expmT2by2
        d = la.GeneralMatrix(np.diag([0.5, -1.0, 2.0]), flags=la.MatrixFlags.Triangular)
        assert _close(d.expm().raw, np.diag(np.exp([0.5, -1.0, 2.0])), 1e-15)
        tri = la.GMat(M4)
        tri.detect(la.MatrixFlags.Triangular)                               # spec/expm_spec.cr:31-32
        assert tri.flags.upper_triangular and _close(tri.expm().raw, M4_EXP, 1e-13)
        with pytest.raises(NotImplementedError):
            la.GMat(M4).expm(schur_fact=True)
        with pytest.raises(ValueError):
            la.GMat(np.ones((2, 3))).expm()
        m = la.GMatComplex([[1.0, 2.0], [-1.0, 3.0]])
        assert _close(la.cosm(m).raw + 1j * la.sinm(m).raw, R.expm(1j * m.raw), 1e-13)
        a = la.GMat([[1.0, 3.0, 0.5], [1.0, 4.0, -1.0], [0.2, 0.1, 0.3]])
        for f in ("cosm", "sinm", "tanm", "coshm", "sinhm", "tanhm"):
            assert _close(getattr(la, f)(a).raw, getattr(R, f)(a.raw), 1e-11), f
