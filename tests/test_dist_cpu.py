"""CPU tests (not gpu) of the multi-rank drivers in linalg_b200/dist.py: world_size = 2 over the
gloo backend, local arithmetic supplied by an oracle-backed ops object, so that the sharding and
exchange logic (block-cyclic ownership, panel broadcast, TSQR all-gather) is exercised without
a GPU.  On the GPU the same drivers run with `DevOps` (libla_b200.so) and NCCL."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from linalg_b200 import dist as ladist


class CpuOps:
    """Oracle-backed local arithmetic on CPU tensors (test infrastructure)."""

    def __init__(self):
        self._info = 0

    def potrf_lower(self, a, k0=0):
        from oracle import c_ref
        F, info = c_ref.potrf("L", np.asfortranarray(a.numpy().T))
        self._info = self._info or (info + k0 if info else 0)
        a.copy_(torch.from_numpy(np.ascontiguousarray(F.T)))

    def trsm_right_lower_ct(self, tri, b):
        import scipy.linalg as sl
        L = np.tril(tri.numpy().T)
        B = b.numpy().T
        X = sl.solve_triangular(L, B.conj().T, lower=True).conj().T
        b.copy_(torch.from_numpy(np.ascontiguousarray(X.T)))

    def gemm_nt_sub(self, c, a, b):
        C, A, B = c.numpy().T, a.numpy().T, b.numpy().T
        c.copy_(torch.from_numpy(np.ascontiguousarray((C - A @ B.conj().T).T)))

    def trsm_left(self, tri, b, *, lower, unit=False, conj_trans=False):
        import scipy.linalg as sl
        Tm = tri.numpy().T
        X = sl.solve_triangular(Tm, b.numpy().T, lower=lower, unit_diagonal=unit,
                                trans="C" if conj_trans else "N")
        b.copy_(torch.from_numpy(np.ascontiguousarray(X.T)))

    def gemm_cn_sub(self, c, a, b):
        C, A, B = c.numpy().T, a.numpy().T, b.numpy().T
        c.copy_(torch.from_numpy(np.ascontiguousarray((C - A.conj().T @ B).T)))

    def gemm_nn_sub(self, c, a, b):
        C, A, B = c.numpy().T, a.numpy().T, b.numpy().T
        c.copy_(torch.from_numpy(np.ascontiguousarray((C - A @ B).T)))

    def gemm_nt_sub_stair(self, c, a, b, nb, row_first, row_step, col_first, col_step):
        # block by block: only the blocks whose global row block >= global column block
        C, A, B = c.numpy().T.copy(), a.numpy().T, b.numpy().T
        for i in range(C.shape[0] // nb):
            for j in range(C.shape[1] // nb):
                if row_first + i * row_step >= col_first + j * col_step:
                    C[i * nb:(i + 1) * nb, j * nb:(j + 1) * nb] -= \
                        A[i * nb:(i + 1) * nb] @ B[j * nb:(j + 1) * nb].conj().T
        c.copy_(torch.from_numpy(np.ascontiguousarray(C.T)))

    def trsm_left_lower_unit(self, tri, b):
        import scipy.linalg as sl
        L = np.tril(tri.numpy().T, -1) + np.eye(tri.shape[0])
        X = sl.solve_triangular(L, b.numpy().T, lower=True, unit_diagonal=True)
        b.copy_(torch.from_numpy(np.ascontiguousarray(X.T)))

    def getrf(self, a, ipiv, row0=0):
        from oracle import c_ref
        LU, piv, info = c_ref.getrf(np.asfortranarray(a.numpy().T))
        self._info = self._info or (info + row0 if info else 0)
        a.copy_(torch.from_numpy(np.ascontiguousarray(LU.T)))
        ipiv.copy_(torch.from_numpy(np.asarray(piv, dtype=np.int32)))

    def laswp(self, a, ipiv, k1, k2):
        M = a.numpy().T.copy()
        for k in range(k1, k2):
            pk = int(ipiv[k]) - 1
            if pk != k:
                M[[k, pk], :] = M[[pk, k], :]
        a.copy_(torch.from_numpy(np.ascontiguousarray(M.T)))

    def info_value(self):
        return self._info

    def geqrf(self, a):
        from oracle import c_ref
        QR, tau = c_ref.geqrf(np.asfortranarray(a.numpy().T))
        a.copy_(torch.from_numpy(np.ascontiguousarray(QR.T)))
        return torch.from_numpy(tau)

    def trtrs_upper(self, r, b):
        import scipy.linalg as sl
        X = sl.solve_triangular(np.triu(r.numpy().T), b.numpy().T, lower=False)
        b.copy_(torch.from_numpy(np.ascontiguousarray(X.T)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _spd(n):
    rng = np.random.default_rng(7)
    G = rng.random((n, n)) - 0.5
    return G @ G.T / n + np.eye(n)


def _chol_worker(rank, world, port, P, Q, n, nb, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        S = _spd(n)
        grid = ladist.Grid(P, Q)
        lay = ladist.BlockCyclic(n, nb, grid)
        rows, cols = lay.global_rows().numpy(), lay.global_cols().numpy()
        a_local = torch.from_numpy(np.ascontiguousarray(S[np.ix_(rows, cols)].T))   # (nloc, mloc)
        info = ladist.potrf_block_cyclic(a_local, n, nb, CpuOps(), grid)
        rhs = np.random.default_rng(5).random((n, 3))
        x = ladist.potrs_block_cyclic(a_local, torch.from_numpy(np.ascontiguousarray(rhs.T)), n, nb,
                                      CpuOps(), grid)
        out[rank] = (info, rows, cols, a_local.numpy().T.copy(), x.numpy().T.copy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("P,Q,n,nb", [(1, 2, 96, 16), (2, 2, 128, 16), (1, 4, 160, 32), (2, 4, 192, 8),
                                      (2, 2, 48, 16)])
def test_block_cyclic_cholesky_on_process_grids(P, Q, n, nb):
    world = P * Q
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_chol_worker, args=(world, _free_port(), P, Q, n, nb, out), nprocs=world, join=True)
    S = _spd(n)
    Lref = np.linalg.cholesky(S)
    L = np.zeros((n, n))
    rhs = np.random.default_rng(5).random((n, 3))
    xref = np.linalg.solve(S, rhs)
    for rank in range(world):
        info, rows, cols, loc, x = out[rank]
        assert info == 0
        L[np.ix_(rows, cols)] = loc
        assert np.abs(x - xref).max() < 1e-11 * np.abs(xref).max()      # cho_solve on every rank
    assert np.allclose(np.tril(L), Lref, atol=1e-12)


def _chol_bad_worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        n, nb = 64, 16
        S = np.eye(n)
        S[37, 37] = -1.0
        grid = ladist.Grid(1, 2)
        lay = ladist.BlockCyclic(n, nb, grid)
        a_local = torch.from_numpy(np.ascontiguousarray(
            S[np.ix_(lay.global_rows().numpy(), lay.global_cols().numpy())].T))
        out[rank] = ladist.potrf_block_cyclic(a_local, n, nb, CpuOps(), grid)
    finally:
        dist.destroy_process_group()


def test_block_cyclic_cholesky_reports_first_bad_pivot_on_every_rank():
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_chol_bad_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    assert out[0] == 38 and out[1] == 38


def _lu_worker(rank, world, port, n, nb, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        A = np.random.default_rng(11).random((n, n))
        grid = ladist.Grid(1, world)
        lay = ladist.BlockCyclic(n, nb, grid)
        cols = lay.global_cols().numpy()
        a_local = torch.from_numpy(np.ascontiguousarray(A[:, cols].T))              # (nloc, n)
        ipiv = torch.zeros(n, dtype=torch.int32)
        info = ladist.getrf_block_cyclic(a_local, ipiv, n, nb, CpuOps(), grid)
        rhs = np.random.default_rng(6).random((n, 2))
        xs = ladist.getrs_block_cyclic(a_local, ipiv, torch.from_numpy(np.ascontiguousarray(rhs.T)),
                                       n, nb, CpuOps(), grid)
        # tall-skinny least squares: rows split over the ranks
        m, nn = 240, 12
        rng = np.random.default_rng(7)
        At = rng.random((m, nn))
        b = rng.random((m, 2))
        r0, r1 = ladist.shard_columns(m, world, rank)
        x = ladist.lstsq_tsqr(torch.from_numpy(np.ascontiguousarray(At[r0:r1].T)),
                              torch.from_numpy(np.ascontiguousarray(b[r0:r1].T)), CpuOps())
        out[rank] = (info, cols, a_local.numpy().T.copy(), ipiv.numpy().copy(), x.numpy().T.copy(),
                     xs.numpy().T.copy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n,nb", [(2, 96, 16), (3, 144, 16), (2, 64, 32)])
def test_block_cyclic_lu_has_lapack_pivots_and_tsqr_matches(world, n, nb):
    from oracle import checks, openblas_ref  # noqa: F401
    import scipy.linalg as sl
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_lu_worker, args=(world, _free_port(), n, nb, out), nprocs=world, join=True)
    A = np.random.default_rng(11).random((n, n))
    LUref, pivref = sl.lu_factor(A)                  # LAPACK dgetrf, 0-based pivots
    LU = np.zeros((n, n))
    xsol = np.linalg.solve(A, np.random.default_rng(6).random((n, 2)))
    for rank in range(world):
        info, cols, loc, ipiv, _, xs = out[rank]
        assert np.abs(xs - xsol).max() < 1e-9 * np.abs(xsol).max()      # LUMatrix#solve
        assert info == 0
        assert np.array_equal(ipiv - 1, pivref)
        LU[:, cols] = loc
    assert np.allclose(LU, LUref, atol=1e-12)
    assert checks.lu_residual(A, LU, out[0][3]) < 1e-12
    m, nn = 240, 12
    rng = np.random.default_rng(7)
    At = rng.random((m, nn))
    b = rng.random((m, 2))
    xref = np.linalg.lstsq(At, b, rcond=None)[0]
    for rank in range(world):
        assert np.allclose(out[rank][4], xref, atol=1e-10)


def test_grid_shapes_and_block_cyclic_index_arithmetic():
    assert [ladist.default_grid_shape(w) for w in (1, 2, 4, 8, 16)] == [(1, 1), (1, 2), (2, 2), (2, 4),
                                                                        (4, 4)]

    class G:                       # a grid position without a process group
        def __init__(self, P, Q, p, q):
            self.P, self.Q, self.p, self.q = P, Q, p, q
    n, nb = 96, 8
    seen = np.zeros((n, n), dtype=int)
    for p in range(2):
        for q in range(4):
            lay = ladist.BlockCyclic(n, nb, G(2, 4, p, q))
            seen[np.ix_(lay.global_rows().numpy(), lay.global_cols().numpy())] += 1
            for g in range(lay.nblk + 1):
                r = lay.lrow_from(g)
                assert all(b < g for b in lay.row_blocks[: r // nb])
                assert all(b >= g for b in lay.row_blocks[r // nb:])
    assert (seen == 1).all()       # every element lives on exactly one rank


def test_shard_columns_partition():
    for n in (1, 7, 16384, 65536):
        for world in (1, 2, 3, 8):
            edges = [ladist.shard_columns(n, world, r) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(edges[i][1] == edges[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in edges]
            assert max(sizes) - min(sizes) <= 1
