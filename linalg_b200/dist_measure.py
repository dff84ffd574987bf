"""Synthetic operands, timing and distributed residual checks for the multi-GPU drivers
(used by bench.py --gpus N and tools/dist_gpu_check.py).

No rank ever holds a whole matrix: every rank builds its own blocks of a matrix defined by a
closed formula of the global indices, and the residuals are formed from distributed
matrix-vector products with the same formula.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist

from . import dist as ladist

CHUNK_COLS = 2048


def _world():
    return dist.get_world_size() if dist.is_initialized() else 1


def _allreduce(t, collective=True):
    if collective and _world() > 1:
        dist.all_reduce(t)
    return t


def spd_block(rows_g: torch.Tensor, cols_g: torch.Tensor, n: int) -> torch.Tensor:
    """S[rows, cols] of the SPD matrix S = n*I + 0.5*sin(12.9898*min(i,j) + 78.233*max(i,j) + 0.1)
    (symmetric by construction, strictly diagonally dominant), as a (len(cols), len(rows)) tensor
    = column-major len(rows) x len(cols)."""
    i = rows_g.to(torch.float64).view(1, -1)
    j = cols_g.to(torch.float64).view(-1, 1)
    s = 0.5 * torch.sin(torch.minimum(i, j) * 12.9898 + torch.maximum(i, j) * 78.233 + 0.1)
    s += float(n) * (i == j)
    return s


def dense_block(rows_g: torch.Tensor, cols_g: torch.Tensor) -> torch.Tensor:
    """A[rows, cols] of a dense pseudo-random matrix with entries in [0, 1) (the shape of
    Mat.rand, src/matrix/matrix.cr:490-492), same layout as spd_block."""
    i = rows_g.to(torch.float64).view(1, -1)
    j = cols_g.to(torch.float64).view(-1, 1)
    return torch.frac(torch.sin(i * 12.9898 + j * 78.233 + 1.0) * 43758.5453).abs()


def _fill(a_local, rows_g, cols_g, gen):
    for c0 in range(0, a_local.shape[0], CHUNK_COLS):
        c1 = min(c0 + CHUNK_COLS, a_local.shape[0])
        a_local[c0:c1] = gen(rows_g, cols_g[c0:c1])


def _timed(fn, barrier=True):
    if barrier and _world() > 1:
        torch.cuda.synchronize()
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if _world() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()), out


def _probe(n, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.rand(n, dtype=torch.float64, generator=g) - 0.5).cuda()


def potrf_residual(a_local, lay, n, collective=True):
    """||S x - L (L^T x)|| / (||S||_F ||x||) for the distributed factor in a_local."""
    rows_g, cols_g = lay.global_rows().cuda(), lay.global_cols().cuda()
    x = _probe(n, 3)
    xr = x[rows_g]
    y = torch.zeros(n, dtype=torch.float64, device="cuda")        # L^T x
    sx = torch.zeros(n, dtype=torch.float64, device="cuda")
    s2 = torch.zeros(1, dtype=torch.float64, device="cuda")
    for c0 in range(0, a_local.shape[0], CHUNK_COLS):
        c1 = min(c0 + CHUNK_COLS, a_local.shape[0])
        mask = rows_g.view(1, -1) >= cols_g[c0:c1].view(-1, 1)
        y[cols_g[c0:c1]] += torch.where(mask, a_local[c0:c1], 0.0) @ xr
        s = spd_block(rows_g, cols_g[c0:c1], n)
        sx.index_add_(0, rows_g, s.T @ x[cols_g[c0:c1]])
        s2 += (s * s).sum()
    _allreduce(y, collective)
    _allreduce(sx, collective)
    _allreduce(s2, collective)
    z = torch.zeros(n, dtype=torch.float64, device="cuda")        # L y
    for c0 in range(0, a_local.shape[0], CHUNK_COLS):
        c1 = min(c0 + CHUNK_COLS, a_local.shape[0])
        mask = rows_g.view(1, -1) >= cols_g[c0:c1].view(-1, 1)
        z.index_add_(0, rows_g, torch.where(mask, a_local[c0:c1], 0.0).T @ y[cols_g[c0:c1]])
    _allreduce(z, collective)
    return float(((sx - z).norm() / (s2.sqrt() * x.norm())).item())


def getrf_residual(a_local, ipiv, lay, n, collective=True):
    """||P A x - L (U x)|| / (||A||_F ||x||) for block columns dealt cyclically (1 x Q grid)."""
    rows_g, cols_g = lay.global_rows().cuda(), lay.global_cols().cuda()   # rows_g = 0..n-1
    x = _probe(n, 4)
    ux = torch.zeros(n, dtype=torch.float64, device="cuda")
    ax = torch.zeros(n, dtype=torch.float64, device="cuda")
    a2 = torch.zeros(1, dtype=torch.float64, device="cuda")
    for c0 in range(0, a_local.shape[0], CHUNK_COLS):
        c1 = min(c0 + CHUNK_COLS, a_local.shape[0])
        upper = rows_g.view(1, -1) <= cols_g[c0:c1].view(-1, 1)
        ux += torch.where(upper, a_local[c0:c1], 0.0).T @ x[cols_g[c0:c1]]
        a = dense_block(rows_g, cols_g[c0:c1])
        ax += a.T @ x[cols_g[c0:c1]]
        a2 += (a * a).sum()
    _allreduce(ux, collective)
    _allreduce(ax, collective)
    _allreduce(a2, collective)
    lux = torch.zeros(n, dtype=torch.float64, device="cuda")
    for c0 in range(0, a_local.shape[0], CHUNK_COLS):
        c1 = min(c0 + CHUNK_COLS, a_local.shape[0])
        lower = rows_g.view(1, -1) > cols_g[c0:c1].view(-1, 1)
        lux += torch.where(lower, a_local[c0:c1], 0.0).T @ ux[cols_g[c0:c1]]
    _allreduce(lux, collective)
    lux += ux                                                       # unit diagonal of L
    piv = ipiv.cpu().numpy().astype(np.int64) - 1
    perm = np.arange(n)
    for i in range(n):
        j = piv[i]
        if j != i:
            perm[i], perm[j] = perm[j], perm[i]
    pax = ax[torch.from_numpy(perm).cuda()]
    return float(((pax - lux).norm() / (a2.sqrt() * x.norm())).item())


def _matvec(gen, lay, n, x, collective=True):
    """A x for the matrix defined by `gen` from this rank's blocks (all-reduced)."""
    rows_g, cols_g = lay.global_rows().cuda(), lay.global_cols().cuda()
    out = torch.zeros(n, dtype=torch.float64, device="cuda")
    for c0 in range(0, len(cols_g), CHUNK_COLS):
        c1 = min(c0 + CHUNK_COLS, len(cols_g))
        out.index_add_(0, rows_g, gen(rows_g, cols_g[c0:c1]).T @ x[cols_g[c0:c1]])
    return _allreduce(out, collective)


def run_potrf(n, nb, grid, reserve_sms=0, reps=1, warm_n=None):
    """Timed distributed Cholesky on the synthetic SPD matrix; returns a dict of results."""
    ops = ladist.DevOps(reserve_sms=reserve_sms)
    if warm_n:                         # first use: NCCL sub-communicators, kernel attributes
        lw = ladist.BlockCyclic(warm_n, nb, grid)
        aw = torch.empty((lw.nloc, lw.mloc), dtype=torch.float64, device="cuda")
        _fill(aw, lw.global_rows().cuda(), lw.global_cols().cuda(), lambda r, c: spd_block(r, c, warm_n))
        ladist.potrf_block_cyclic(aw, warm_n, nb, ops, grid)
        del aw
    lay = ladist.BlockCyclic(n, nb, grid)
    a_local = torch.empty((lay.nloc, lay.mloc), dtype=torch.float64, device="cuda")
    rows_g, cols_g = lay.global_rows().cuda(), lay.global_cols().cuda()
    best, info = 1e30, 0
    for _ in range(reps):
        _fill(a_local, rows_g, cols_g, lambda r, c: spd_block(r, c, n))
        ms, info = _timed(lambda: ladist.potrf_block_cyclic(a_local, n, nb, ops, grid))
        best = min(best, ms)
    res = potrf_residual(a_local, lay, n)
    # cho_solve with the distributed factor: S x = S x0, one right-hand side
    x0 = _probe(n, 9)
    rhs = _matvec(lambda r, c: spd_block(r, c, n), lay, n, x0).view(1, n)
    solve_ms, xs = _timed(lambda: ladist.potrs_block_cyclic(a_local, rhs, n, nb, ops, grid))
    solve_err = float(((xs[0] - x0).abs().max() / x0.abs().max()).item())
    del a_local
    torch.cuda.empty_cache()
    return {"n": n, "nb": nb, "grid": f"{grid.P}x{grid.Q}", "ms": best, "info": info,
            "tflops": n ** 3 / 3.0 / best * 1e-9, "residual": res, "reserve_sms": reserve_sms,
            "solve_ms": solve_ms, "solve_max_rel_err": solve_err}


def run_getrf(n, nb, grid, reserve_sms=0, reps=1, warm_n=None):
    """Timed distributed LU on the synthetic dense matrix (block columns over a 1 x Q grid)."""
    ops = ladist.DevOps(reserve_sms=reserve_sms)
    if warm_n:
        lw = ladist.BlockCyclic(warm_n, nb, grid)
        aw = torch.empty((lw.nloc, warm_n), dtype=torch.float64, device="cuda")
        _fill(aw, lw.global_rows().cuda(), lw.global_cols().cuda(), dense_block)
        ladist.getrf_block_cyclic(aw, torch.zeros(warm_n, dtype=torch.int32, device="cuda"), warm_n,
                                  nb, ops, grid)
        del aw
    lay = ladist.BlockCyclic(n, nb, grid)
    a_local = torch.empty((lay.nloc, n), dtype=torch.float64, device="cuda")
    ipiv = torch.zeros(n, dtype=torch.int32, device="cuda")
    rows_g, cols_g = lay.global_rows().cuda(), lay.global_cols().cuda()
    best, info = 1e30, 0
    for _ in range(reps):
        _fill(a_local, rows_g, cols_g, dense_block)
        ms, info = _timed(lambda: ladist.getrf_block_cyclic(a_local, ipiv, n, nb, ops, grid))
        best = min(best, ms)
    res = getrf_residual(a_local, ipiv, lay, n)
    # LA.solve with the distributed factors: A x = A x0, one right-hand side; the error is judged
    # through the residual (a random dense matrix of this size is not well conditioned)
    x0 = _probe(n, 10)
    rhs = _matvec(dense_block, lay, n, x0).view(1, n)
    b0 = rhs.clone()
    solve_ms, xs = _timed(lambda: ladist.getrs_block_cyclic(a_local, ipiv, rhs, n, nb, ops, grid))
    # a_local holds the factors now: the residual uses the generator again
    r = _matvec(dense_block, lay, n, xs[0]) - b0[0]
    solve_res = float((r.norm() / (float(n) * 0.5774 * xs[0].norm())).item())   # ||A||_F ~ n / sqrt(3)
    del a_local
    torch.cuda.empty_cache()
    return {"n": n, "nb": nb, "grid": f"1x{grid.Q}", "ms": best, "info": info,
            "tflops": 2.0 * n ** 3 / 3.0 / best * 1e-9, "residual": res, "reserve_sms": reserve_sms,
            "solve_ms": solve_ms, "solve_residual": solve_res}


class _OneRank:
    """Grid position of a single-GPU run (no process group involved)."""
    P = Q = 1
    p = q = 0
    world, rank = 1, 0
    row_group = col_group = None


def run_single_gpu(kind, n):
    """The single-GPU library routine (dpotrf / dgetrf device API) on the same synthetic matrix:
    the denominator of the multi-GPU speed-up.  Runs on the calling rank only (no collective)."""
    from . import dev
    lay = ladist.BlockCyclic(n, 512 if n % 512 == 0 else n, _OneRank)
    a = torch.empty((n, n), dtype=torch.float64, device="cuda")
    rows_g = torch.arange(n, device="cuda")
    info = torch.zeros(1, dtype=torch.int32, device="cuda")
    ipiv = torch.zeros(n, dtype=torch.int32, device="cuda")
    gen = (lambda r, c: spd_block(r, c, n)) if kind == "potrf" else dense_block
    best = 1e30
    for _ in range(2):
        _fill(a, rows_g, rows_g, gen)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        if kind == "potrf":
            dev.potrf(dev.LOWER, n, a, n, info)
        else:
            dev.getrf(n, n, a, n, ipiv, info)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    res = (potrf_residual(a, lay, n, collective=False) if kind == "potrf"
           else getrf_residual(a, ipiv, lay, n, collective=False))
    flops = n ** 3 / 3.0 if kind == "potrf" else 2.0 * n ** 3 / 3.0
    out = {"n": n, "ms": best, "info": int(info.item()), "tflops": flops / best * 1e-9, "residual": res}
    del a
    torch.cuda.empty_cache()
    return out


def run_tsqr(m_total, n, reps=2):
    """Timed TSQR least squares: m_total x n, rows split evenly over the ranks, consistent
    right-hand side (exact solution known)."""
    world, rank = _world(), (dist.get_rank() if dist.is_initialized() else 0)
    m_loc = m_total // world
    g = torch.Generator(device="cuda").manual_seed(11 + rank)
    A = torch.rand((n, m_loc), dtype=torch.float64, device="cuda", generator=g)
    xt = torch.rand((1, n), dtype=torch.float64, device="cuda",
                    generator=torch.Generator(device="cuda").manual_seed(5))
    b = xt @ A
    ops = ladist.DevOps()
    best, xs = 1e30, None
    for _ in range(reps + 1):
        a1, b1 = A.clone(), b.clone()
        ms, xs = _timed(lambda: ladist.lstsq_tsqr(a1, b1, ops))
        best = min(best, ms)
    err = float(((xs - xt).abs().max() / xt.abs().max()).item())
    return {"m": m_total, "n": n, "ms": best, "max_rel_err": err,
            "tflops": (2.0 * m_total * n * n - 2.0 / 3.0 * n ** 3) / best * 1e-9}
