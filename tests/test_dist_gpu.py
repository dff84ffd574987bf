"""GPU tests of linalg_b200/dist.py with the product `DevOps` (libla_b200.so device API) on one
rank, and of the device entry points only the multi-GPU drivers use (staircase-masked product,
row interchanges).  The multi-rank exchange logic is covered on CPU by tests/test_dist_cpu.py
(gloo, process grids up to 2 x 4) and on real GPUs by `bench.py --gpus N`, which runs the
drivers under torchrun and reports their residuals."""
import numpy as np
import pytest
import torch

from linalg_b200 import dev
from linalg_b200 import dist as ladist

pytestmark = pytest.mark.gpu


def test_block_cyclic_cholesky_single_rank_matches_oracle(oracle_abi):
    n, nb = 640, 128
    rng = np.random.default_rng(3)
    G = rng.random((n, n)) - 0.5
    S = G @ G.T / n + np.eye(n)
    a = torch.from_numpy(np.ascontiguousarray(S.T)).cuda()
    info = ladist.potrf_block_cyclic(a, n, nb, ladist.DevOps())
    assert info == 0
    L = np.tril(a.cpu().numpy().T)
    assert np.abs(L @ L.T - S).max() / np.abs(S).max() < 1e-12
    assert np.allclose(L, np.linalg.cholesky(S), atol=1e-11)
    rhs = rng.random((n, 3))
    x = ladist.potrs_block_cyclic(a, torch.from_numpy(np.ascontiguousarray(rhs.T)).cuda(), n, nb,
                                  ladist.DevOps())
    assert np.abs(x.cpu().numpy().T - np.linalg.solve(S, rhs)).max() < 1e-11


def test_block_cyclic_cholesky_reports_first_bad_pivot():
    n, nb = 256, 64
    S = np.eye(n)
    S[100, 100] = -1.0
    a = torch.from_numpy(np.ascontiguousarray(S.T)).cuda()
    assert ladist.potrf_block_cyclic(a, n, nb, ladist.DevOps()) == 101


def test_block_cyclic_lu_single_rank_has_lapack_pivots():
    import scipy.linalg as sl
    from oracle import checks
    n, nb = 768, 128
    A = np.random.default_rng(5).random((n, n))
    a = torch.from_numpy(np.ascontiguousarray(A.T)).cuda()
    ipiv = torch.zeros(n, dtype=torch.int32, device="cuda")
    assert ladist.getrf_block_cyclic(a, ipiv, n, nb, ladist.DevOps()) == 0
    LUref, pivref = sl.lu_factor(A)
    assert np.array_equal(ipiv.cpu().numpy() - 1, pivref)
    assert checks.lu_residual(A, a.cpu().numpy().T, ipiv.cpu().numpy()) < 1e-12
    rhs = np.random.default_rng(6).random((n, 2))
    x = ladist.getrs_block_cyclic(a, ipiv, torch.from_numpy(np.ascontiguousarray(rhs.T)).cuda(), n, nb,
                                  ladist.DevOps())
    assert checks.solve_residual(A, x.cpu().numpy().T, rhs) < 1e-12


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32, torch.complex128])
@pytest.mark.parametrize("m,n,k,nb,rf,rs,cf,cs", [(1024, 512, 128, 128, 3, 2, 2, 4),
                                                  (1536, 1536, 256, 256, 1, 1, 1, 1),
                                                  (640, 384, 64, 128, 0, 2, 5, 4)])
def test_staircase_product_matches_blockwise_reference(dtype, m, n, k, nb, rf, rs, cf, cs):
    """C <- C - A B^H on the blocks whose global row block >= global column block; every other
    block of C must come back bit-identical."""
    g = torch.Generator(device="cuda").manual_seed(9)
    mk = lambda *s: (torch.rand(*s, dtype=dtype, device="cuda", generator=g) - 0.5)  # noqa: E731
    A, B, C0 = mk(k, m), mk(k, n), mk(n, m)
    C = C0.clone()
    dev.gemm_stair(dev.NO_TRANS, dev.CONJ_TRANS, m, n, k, -1.0, A, m, B, n, 1.0, C, m, nb, rf, rs, cf, cs)
    ref = C0.clone()
    hi = torch.complex128 if dtype.is_complex else torch.float64
    full = (C0.T.to(hi) - A.T.to(hi) @ B.T.to(hi).conj().T).to(dtype).T
    for i in range(m // nb):
        for j in range(n // nb):
            if rf + i * rs >= cf + j * cs:
                ref[j * nb:(j + 1) * nb, i * nb:(i + 1) * nb] = full[j * nb:(j + 1) * nb, i * nb:(i + 1) * nb]
    tol = 1e-5 if dtype == torch.float32 else 1e-12
    assert float((C - ref).abs().max()) <= tol * float(ref.abs().max()) * k ** 0.5
    for i in range(m // nb):
        for j in range(n // nb):
            if rf + i * rs < cf + j * cs:
                assert torch.equal(C[j * nb:(j + 1) * nb, i * nb:(i + 1) * nb],
                                   C0[j * nb:(j + 1) * nb, i * nb:(i + 1) * nb])
    if dtype == torch.float64:
        assert dev.last_gemm_path() == 1, "the masked tensor-core launch was not taken"


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32, torch.complex128])
def test_laswp_dev_applies_the_interchanges_in_order(dtype):
    m, n = 3000, 257
    rng = np.random.default_rng(2)
    piv = np.array([rng.integers(k, m) + 1 for k in range(700)], dtype=np.int32)
    g = torch.Generator(device="cuda").manual_seed(1)
    A = torch.rand((n, m), dtype=dtype, device="cuda", generator=g)
    ref = A.cpu().numpy().T.copy()
    for k in range(100, 700):
        p_ = piv[k] - 1
        if p_ != k:
            ref[[k, p_], :] = ref[[p_, k], :]
    dev.laswp(m, n, A, m, 100, 700, torch.from_numpy(piv).cuda())
    assert np.array_equal(A.cpu().numpy().T, ref)


def test_tsqr_single_rank_matches_lstsq():
    m, n, nrhs = 4096, 48, 3
    rng = np.random.default_rng(4)
    A = rng.random((m, n))
    b = rng.random((m, nrhs))
    x = ladist.lstsq_tsqr(torch.from_numpy(np.ascontiguousarray(A.T)).cuda(),
                          torch.from_numpy(np.ascontiguousarray(b.T)).cuda(), ladist.DevOps())
    xref = np.linalg.lstsq(A, b, rcond=None)[0]
    assert np.abs(x.cpu().numpy().T - xref).max() < 1e-10
