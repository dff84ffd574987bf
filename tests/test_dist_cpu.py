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


def _worker(rank, world, port, n, nb, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(7)
        G = rng.random((n, n)) - 0.5
        S = G @ G.T / n + np.eye(n)
        nblk = (n + nb - 1) // nb
        mine = [k for k in range(nblk) if ladist.owner_of_block(k, world) == rank]
        cols = np.concatenate([np.arange(k * nb, min(n, (k + 1) * nb)) for k in mine])
        a_local = torch.from_numpy(np.ascontiguousarray(S[:, cols].T))       # (ncols, n)
        info = ladist.potrf_block_cyclic(a_local, n, nb, CpuOps())
        # tall-skinny least squares: rows split over the ranks
        m, nn = 240, 12
        A = rng.random((m, nn))
        b = rng.random((m, 2))
        r0, r1 = ladist.shard_columns(m, world, rank)
        x = ladist.lstsq_tsqr(torch.from_numpy(np.ascontiguousarray(A[r0:r1].T)),
                              torch.from_numpy(np.ascontiguousarray(b[r0:r1].T)), CpuOps())
        out[rank] = (info, cols, a_local.numpy().T.copy(), x.numpy().T.copy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n,nb", [(96, 16), (100, 32)])
def test_block_cyclic_cholesky_and_tsqr_world2(n, nb):
    world = 2
    port = _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, port, n, nb, out), nprocs=world, join=True)
    rng = np.random.default_rng(7)
    G = rng.random((n, n)) - 0.5
    S = G @ G.T / n + np.eye(n)
    Lref = np.linalg.cholesky(S)
    L = np.zeros((n, n))
    for rank in range(world):
        info, cols, loc, _ = out[rank]
        assert info == 0
        L[:, cols] = loc
    assert np.allclose(np.tril(L), Lref, atol=1e-12)
    m, nn = 240, 12
    A = rng.random((m, nn))
    b = rng.random((m, 2))
    xref = np.linalg.lstsq(A, b, rcond=None)[0]
    for rank in range(world):
        assert np.allclose(out[rank][3], xref, atol=1e-10)


def test_shard_columns_partition():
    for n in (1, 7, 16384, 65536):
        for world in (1, 2, 3, 8):
            edges = [ladist.shard_columns(n, world, r) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(edges[i][1] == edges[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in edges]
            assert max(sizes) - min(sizes) <= 1
