"""Multi-GPU drivers: one process per GPU, `torch.distributed` for the plumbing.

The reference has no distributed path; what shards naturally is stated in SURVEY.md §8(e):

  * GEMM: B and C split by block columns over the ranks, A replicated, no data-path collective
    (`shard_columns`, used by bench.py).
  * Cholesky: block columns dealt round-robin to the ranks (1-D block-cyclic).  Per block column
    k the owner factors the diagonal block and solves the panel below it, the panel is
    broadcast (NCCL over NVLink on GPUs), and every rank applies the rank-nb update to the block
    columns it owns (`potrf_block_cyclic`).  One broadcast of (N - k*nb) x nb elements per step
    is the only exchange.
  * Tall-skinny least squares: rows split over the ranks, local QR of [A_i | b_i], all-gather of
    the small R factors, QR of the stack, triangular solve (`lstsq_tsqr`).

All matrices are column-major and are held as torch tensors in "transposed" orientation: a
tensor `t` of shape (cols, rows) with stride (ld, 1) is the matrix whose element (r, c) is
t[c, r].  The local arithmetic goes through an `ops` object: `DevOps` (libla_b200.so device API,
the product path) or any object with the same methods (the CPU tests pass an oracle-backed one
and the gloo backend, so that the sharding / exchange logic runs without a GPU).
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_columns(n: int, world: int, rank: int):
    """Contiguous block-column range [c0, c1) of rank `rank`."""
    return (n * rank) // world, (n * (rank + 1)) // world


def owner_of_block(k: int, world: int) -> int:
    return k % world


def local_block_index(k: int, world: int) -> int:
    return k // world


class DevOps:
    """Local arithmetic on CUDA tensors through the device-resident C API."""

    def __init__(self):
        from . import dev
        self.dev = dev
        self.info = torch.zeros(1, dtype=torch.int32, device="cuda")
        self.first = torch.zeros(1, dtype=torch.int32, device="cuda")   # first failing column

    @staticmethod
    def _ld(t):
        return t.stride(0) if t.shape[0] > 1 else max(t.shape[1], 1)

    def potrf_lower(self, a, k0=0):                # a: n x n view at global offset k0
        n = a.shape[0]
        self.dev.potrf(self.dev.LOWER, n, a, self._ld(a), self.info)
        # keep LAPACK's meaning of info (first non-positive pivot, global index) without a sync
        self.first = torch.where((self.first == 0) & (self.info != 0), self.info + k0, self.first)

    def trsm_right_lower_ct(self, tri, b):         # b <- b * tri^-H ; b: m x n view, tri: n x n
        m, n = b.shape[1], b.shape[0]
        self.dev.trsm(self.dev.RIGHT, self.dev.LOWER, self.dev.CONJ_TRANS, self.dev.NON_UNIT, m, n,
                      1.0, tri, self._ld(tri), b, self._ld(b))

    def gemm_nt_sub(self, c, a, b):                # c <- c - a * b^H ; a: m x k, b: n x k, c: m x n
        m, n, k = c.shape[1], c.shape[0], a.shape[0]
        self.dev.gemm(self.dev.NO_TRANS, self.dev.CONJ_TRANS, m, n, k, -1.0, a, self._ld(a), b,
                      self._ld(b), 1.0, c, self._ld(c))

    def info_value(self):
        return int(self.first.item())

    def geqrf(self, a):                            # in-place QR of the m x n view, returns tau
        m, n = a.shape[1], a.shape[0]
        tau = torch.zeros(min(m, n), dtype=a.dtype, device=a.device)
        self.dev.geqrf(m, n, a, self._ld(a), tau)
        return tau

    def trtrs_upper(self, r, b):                   # b <- r^-1 b ; r: n x n upper, b: n x nrhs
        n, nrhs = r.shape[0], b.shape[0]
        self.dev.trtrs(self.dev.UPPER, self.dev.NO_TRANS, self.dev.NON_UNIT, n, nrhs, r, self._ld(r),
                       b, self._ld(b), self.info)


def potrf_block_cyclic(a_local: torch.Tensor, n: int, nb: int, ops, group=None):
    """In-place lower Cholesky of an n x n SPD matrix distributed by block columns.

    `a_local` has shape (nloc * nb_cols, n): the block columns k with k % world == rank, in
    increasing k, each nb columns wide (the last one may be narrower), full height n.  On return
    the lower triangle of every owned block column holds L.  Returns LAPACK's info (0 = ok).

    Look-ahead: as soon as panel k has arrived, the owner of block column k+1 updates that column
    alone, factors it and posts its broadcast (asynchronous, on NCCL's stream), and only then do
    all ranks apply panel k to the rest of their columns — the next panel travels while the
    tensor cores are busy, and no rank ever waits for a barrier, only for the panel it needs.
    """
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    nblk = (n + nb - 1) // nb
    # column offset of each owned block column inside a_local
    offs, pos = {}, 0
    for k in range(nblk):
        if owner_of_block(k, world) == rank:
            offs[k] = pos
            pos += min(nb, n - k * nb)
    bufs = [torch.empty(nb * n, dtype=a_local.dtype, device=a_local.device) for _ in range(2)]

    def factor_and_post(k):
        """Owner: factor block column k (all updates applied) and stage it; all: post the
        broadcast.  Returns (panel view (kb, n - k0) with ld = n - k0, pending work or None)."""
        k0 = k * nb
        kb = min(nb, n - k0)
        owner = owner_of_block(k, world)
        panel = bufs[k % 2][: kb * (n - k0)].view(kb, n - k0)
        if rank == owner:
            col = a_local[offs[k]: offs[k] + kb]       # n x kb view
            diag = col[:, k0: k0 + kb]
            ops.potrf_lower(diag, k0)
            if n - k0 - kb > 0:
                ops.trsm_right_lower_ct(diag, col[:, k0 + kb:])
            panel.copy_(col[:, k0:])
        work = None
        if world > 1:
            work = dist.broadcast(panel, src=owner, group=group, async_op=True)
        return panel, work

    def update(j, k, panel):
        """Block column j (owned) -= panel_k(j0:, :) * panel_k(j0:j0+jb, :)^H."""
        k0, j0 = k * nb, j * nb
        jb = min(nb, n - j0)
        c = a_local[offs[j]: offs[j] + jb][:, j0:]                 # (n - j0) x jb
        ops.gemm_nt_sub(c, panel[:, j0 - k0:], panel[:, j0 - k0: j0 - k0 + jb])

    panel, work = factor_and_post(0)
    for k in range(nblk):
        if work is not None:
            work.wait()
        nxt = (None, None)
        if k + 1 < nblk:
            if owner_of_block(k + 1, world) == rank:
                update(k + 1, k, panel)
            nxt = factor_and_post(k + 1)
        for j in range(k + 2, nblk):
            if owner_of_block(j, world) == rank:
                update(j, k, panel)
        panel, work = nxt
    info = ops.info_value()
    if world > 1:                                   # smallest non-zero info over the ranks
        big = 2 ** 31 - 1
        t = torch.tensor([info if info > 0 else big], dtype=torch.int64, device=a_local.device)
        dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
        info = int(t.item())
        info = 0 if info == big else info
    return info


def lstsq_tsqr(a_local: torch.Tensor, b_local: torch.Tensor, ops, group=None):
    """min ||A x - b|| for a tall matrix split by rows over the ranks (TSQR).

    a_local: (n, m_local) = the m_local x n row block of A; b_local: (nrhs, m_local).
    Returns x as a (nrhs, n) tensor (every rank gets the same copy).  a_local / b_local are
    overwritten.
    """
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    n, nrhs = a_local.shape[0], b_local.shape[0]
    m_local = a_local.shape[1]
    assert m_local >= n + nrhs, "each rank needs at least n + nrhs rows"
    # local QR of [A_i | b_i]: its R factor carries R_i and Q_i^H b_i (and the residual norm)
    ab = torch.cat([a_local, b_local], dim=0).contiguous()          # (n + nrhs) x m_local
    ops.geqrf(ab)
    w = n + nrhs
    r_loc = torch.triu(ab[:, :w].T).T.contiguous()                  # w x w upper triangle
    if world > 1:
        gathered = [torch.empty_like(r_loc) for _ in range(world)]
        dist.all_gather(gathered, r_loc, group=group)
        stack = torch.cat(gathered, dim=1).contiguous()             # (world * w) x w
        ops.geqrf(stack)
        r_all = torch.triu(stack[:, :w].T).T.contiguous()
    else:
        r_all = r_loc
    r = r_all[:n, :n].contiguous()                                  # n x n upper
    c = r_all[n:, :n].contiguous()                                  # nrhs x n  (Q^H b)(0:n)
    ops.trtrs_upper(r, c)
    return c
