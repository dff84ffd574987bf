"""GPU tests of linalg_b200/dist.py with the product `DevOps` (libla_b200.so device API) on one
rank; the world_size-2 exchange logic is covered on CPU by tests/test_dist_cpu.py and on two
GPUs by tools/dist_gpu_check.py (torchrun)."""
import numpy as np
import pytest
import torch

from linalg_b200 import dist as ladist

pytestmark = pytest.mark.gpu


def test_block_column_cholesky_single_rank_matches_oracle(oracle_abi):
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


def test_block_column_cholesky_reports_first_bad_pivot():
    n, nb = 256, 64
    S = np.eye(n)
    S[100, 100] = -1.0
    a = torch.from_numpy(np.ascontiguousarray(S.T)).cuda()
    assert ladist.potrf_block_cyclic(a, n, nb, ladist.DevOps()) == 101


def test_tsqr_single_rank_matches_lstsq():
    m, n, nrhs = 4096, 48, 3
    rng = np.random.default_rng(4)
    A = rng.random((m, n))
    b = rng.random((m, nrhs))
    x = ladist.lstsq_tsqr(torch.from_numpy(np.ascontiguousarray(A.T)).cuda(),
                          torch.from_numpy(np.ascontiguousarray(b.T)).cuda(), ladist.DevOps())
    xref = np.linalg.lstsq(A, b, rcond=None)[0]
    assert np.abs(x.cpu().numpy().T - xref).max() < 1e-10
