"""Multi-GPU drivers: one process per GPU, `torch.distributed` for the plumbing.

The reference has no distributed path; what shards naturally is stated in SURVEY.md §8(e):

  * GEMM: B and C split by block columns over the ranks, A replicated, no data-path collective
    (`shard_columns`, used by bench.py); with operands in host memory every rank uploads 1/N of A
    and the slices are all-gathered over NVLink (`HostGemmPlan`).
  * Cholesky (`potrf_block_cyclic`): 2-D block-cyclic layout on a P x Q process grid (block
    (i, j) lives on grid position (i mod P, j mod Q)).  Per block column k: the owner of the
    diagonal block factors it and sends it down its grid column; the P ranks of that column solve
    their rows of the panel; every rank then receives exactly the two pieces of the panel it
    needs — the rows of its own row blocks (broadcast along its grid row) and the rows of its own
    column blocks (broadcast along its grid column) — and applies ONE rank-nb update to its local
    trailing matrix, restricted to the blocks on or below the diagonal (staircase-masked product).
    Look-ahead: the ranks that hold block column k+1 update that column first, factor it and
    start its exchange on a second stream before they touch the rest of the trailing matrix.
    `potrs_block_cyclic` solves with the distributed factor (cho_solve).
  * LU (`getrf_block_cyclic`): block columns dealt cyclically to the ranks (a 1 x Q grid), so a
    panel and its pivot search stay on one GPU; the factored panel and its pivots are broadcast,
    every rank applies the interchanges, the triangular solve and the rank-nb update to its own
    columns, with the same look-ahead.  `getrs_block_cyclic` solves with the distributed factors.
  * Tall-skinny least squares (`lstsq_tsqr`): rows split over the ranks, local QR of [A_i | b_i],
    all-gather of the small R factors, QR of the stack, triangular solve.

All matrices are column-major and are held as torch tensors in "transposed" orientation: a
tensor `t` of shape (cols, rows) with stride (ld, 1) is the matrix whose element (r, c) is
t[c, r].  The local arithmetic goes through an `ops` object: `DevOps` (libla_b200.so device API,
the product path) or any object with the same methods (the CPU tests pass an oracle-backed one
and the gloo backend, so that the sharding / exchange logic runs without a GPU).
"""
from __future__ import annotations

import contextlib

import torch
import torch.distributed as dist


def shard_columns(n: int, world: int, rank: int):
    """Contiguous block-column range [c0, c1) of rank `rank`."""
    return (n * rank) // world, (n * (rank + 1)) // world


def default_grid_shape(world: int):
    """P x Q with P <= Q, P | Q: 1x1, 1x2, 2x2, 2x4 for 1, 2, 4, 8 ranks."""
    p = 1
    while (p * 2) * (p * 2) <= world and world % (p * 2) == 0 and (world // (p * 2)) % (p * 2) == 0:
        p *= 2
    return p, world // p


class Grid:
    """P x Q process grid over the ranks of `torch.distributed`'s world (row-major: rank = p*Q+q)
    with one sub-group per grid row and per grid column.  Every rank must construct it (group
    creation is collective).  Requires P | Q, so that the rows of a panel that grid column q needs
    all sit in ONE grid row (q mod P)."""

    def __init__(self, P: int | None = None, Q: int | None = None, comm_ctas: int = 0):
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        if P is None or Q is None:
            P, Q = default_grid_shape(self.world)
        if P * Q != self.world or Q % P != 0:
            raise ValueError(f"grid {P} x {Q} does not fit {self.world} ranks (need P*Q = world, P | Q)")
        self.P, self.Q = P, Q
        self.p, self.q = self.rank // Q, self.rank % Q
        self.row_group = self.col_group = None
        if self.world > 1:
            opts = None
            if comm_ctas > 0 and dist.get_backend() == "nccl":
                # keep the collectives on a few SMs: they run beside the trailing update
                opts = dist.ProcessGroupNCCL.Options()
                opts.config.max_ctas = comm_ctas
                opts.config.min_ctas = 1
            for pp in range(P):
                g = dist.new_group([pp * Q + qq for qq in range(Q)], pg_options=opts) if Q > 1 else None
                if pp == self.p:
                    self.row_group = g
            for qq in range(Q):
                g = dist.new_group([pp * Q + qq for pp in range(P)], pg_options=opts) if P > 1 else None
                if qq == self.q:
                    self.col_group = g

    def rank_of(self, p: int, q: int) -> int:
        return p * self.Q + q


class BlockCyclic:
    """Index arithmetic of an n x n matrix in nb x nb blocks on a Grid (n must be a multiple of
    nb).  The local matrix of grid position (p, q) holds the row blocks p, p+P, ... and the column
    blocks q, q+Q, ... in increasing order: local shape mloc x nloc, column-major."""

    def __init__(self, n: int, nb: int, grid: Grid):
        if n % nb != 0:
            raise ValueError("block-cyclic drivers need n to be a multiple of nb")
        self.n, self.nb, self.grid = n, nb, grid
        self.nblk = n // nb
        self.row_blocks = list(range(grid.p, self.nblk, grid.P))
        self.col_blocks = list(range(grid.q, self.nblk, grid.Q))
        self.mloc = len(self.row_blocks) * nb
        self.nloc = len(self.col_blocks) * nb

    def lrow_from(self, g: int) -> int:
        """Local row offset of the first owned row block with global index >= g."""
        li = max(0, -(-(g - self.grid.p) // self.grid.P))
        return min(li * self.nb, self.mloc)

    def lcol_from(self, g: int) -> int:
        lj = max(0, -(-(g - self.grid.q) // self.grid.Q))
        return min(lj * self.nb, self.nloc)

    def global_rows(self) -> torch.Tensor:
        """Global row index of every local row (int64, CPU)."""
        return torch.cat([torch.arange(i * self.nb, (i + 1) * self.nb) for i in self.row_blocks]) \
            if self.row_blocks else torch.zeros(0, dtype=torch.int64)

    def global_cols(self) -> torch.Tensor:
        return torch.cat([torch.arange(j * self.nb, (j + 1) * self.nb) for j in self.col_blocks]) \
            if self.col_blocks else torch.zeros(0, dtype=torch.int64)


class _SideStream:
    """The exchange runs on a second CUDA stream next to the compute stream; on CPU tensors
    (gloo tests) every method is a no-op and the collectives simply block."""

    def __init__(self, device: torch.device):
        self.cuda = device.type == "cuda"
        self.stream = torch.cuda.Stream(device) if self.cuda else None

    def section(self):
        """`with side.section():` work issued inside is ordered after everything already on the
        compute stream and runs on the side stream."""
        if not self.cuda:
            return contextlib.nullcontext()
        self.stream.wait_stream(torch.cuda.current_stream())
        return torch.cuda.stream(self.stream)

    def mark(self):
        """Inside a section: an event the compute stream can wait for."""
        if not self.cuda:
            return None
        ev = torch.cuda.Event()
        ev.record(self.stream)
        return ev

    def wait(self, ev):
        if ev is not None:
            torch.cuda.current_stream().wait_event(ev)


class DevOps:
    """Local arithmetic on CUDA tensors through the device-resident C API.

    `reserve_sms` > 0: the long products of the trailing updates leave that many SMs free, so
    that the collectives of the NEXT panel (kernels on another stream) can run beside them.
    """

    def __init__(self, reserve_sms: int = 0):
        from . import dev
        self.dev = dev
        self.info = torch.zeros(1, dtype=torch.int32, device="cuda")
        self.first = torch.zeros(1, dtype=torch.int32, device="cuda")   # first failing column
        self.reserve_sms = reserve_sms
        self.sm_count = torch.cuda.get_device_properties(torch.cuda.current_device()).multi_processor_count

    @staticmethod
    def _ld(t):
        return t.stride(0) if t.shape[0] > 1 else max(t.shape[1], 1)

    @contextlib.contextmanager
    def _limited(self):
        if self.reserve_sms > 0:
            old = self.dev.set_sm_limit(self.sm_count - self.reserve_sms)
            try:
                yield
            finally:
                self.dev.set_sm_limit(old)
        else:
            yield

    def potrf_lower(self, a, k0=0):                # a: n x n view at global offset k0
        n = a.shape[0]
        self.dev.potrf(self.dev.LOWER, n, a, self._ld(a), self.info)
        # keep LAPACK's meaning of info (first non-positive pivot, global index) without a sync
        self.first = torch.where((self.first == 0) & (self.info != 0), self.info + k0, self.first)

    def trsm_right_lower_ct(self, tri, b):         # b <- b * tri^-H ; b: m x n view, tri: n x n
        m, n = b.shape[1], b.shape[0]
        self.dev.trsm(self.dev.RIGHT, self.dev.LOWER, self.dev.CONJ_TRANS, self.dev.NON_UNIT, m, n,
                      1.0, tri, self._ld(tri), b, self._ld(b))

    def trsm_left_lower_unit(self, tri, b):        # b <- tri^-1 b ; tri: m x m unit lower, b: m x n
        m, n = b.shape[1], b.shape[0]
        self.dev.trsm(self.dev.LEFT, self.dev.LOWER, self.dev.NO_TRANS, self.dev.UNIT, m, n, 1.0,
                      tri, self._ld(tri), b, self._ld(b))

    def trsm_left(self, tri, b, *, lower, unit=False, conj_trans=False):
        """b <- op(tri)^-1 b with tri the named triangle of an m x m block, b: m x n."""
        m, n = b.shape[1], b.shape[0]
        d = self.dev
        d.trsm(d.LEFT, d.LOWER if lower else d.UPPER, d.CONJ_TRANS if conj_trans else d.NO_TRANS,
               d.UNIT if unit else d.NON_UNIT, m, n, 1.0, tri, self._ld(tri), b, self._ld(b))

    def gemm_cn_sub(self, c, a, b):                # c <- c - a^H * b ; a: k x m, b: k x n, c: m x n
        m, n, k = c.shape[1], c.shape[0], a.shape[1]
        self.dev.gemm(self.dev.CONJ_TRANS, self.dev.NO_TRANS, m, n, k, -1.0, a, self._ld(a), b,
                      self._ld(b), 1.0, c, self._ld(c))

    def gemm_nt_sub(self, c, a, b):                # c <- c - a * b^H ; a: m x k, b: n x k, c: m x n
        m, n, k = c.shape[1], c.shape[0], a.shape[0]
        with self._limited():
            self.dev.gemm(self.dev.NO_TRANS, self.dev.CONJ_TRANS, m, n, k, -1.0, a, self._ld(a), b,
                          self._ld(b), 1.0, c, self._ld(c))

    def gemm_nn_sub(self, c, a, b):                # c <- c - a * b ; a: m x k, b: k x n, c: m x n
        m, n, k = c.shape[1], c.shape[0], a.shape[0]
        with self._limited():
            self.dev.gemm(self.dev.NO_TRANS, self.dev.NO_TRANS, m, n, k, -1.0, a, self._ld(a), b,
                          self._ld(b), 1.0, c, self._ld(c))

    def gemm_nt_sub_stair(self, c, a, b, nb, row_first, row_step, col_first, col_step):
        """c <- c - a * b^H on the blocks whose global row block >= global column block."""
        m, n, k = c.shape[1], c.shape[0], a.shape[0]
        with self._limited():
            self.dev.gemm_stair(self.dev.NO_TRANS, self.dev.CONJ_TRANS, m, n, k, -1.0, a, self._ld(a),
                                b, self._ld(b), 1.0, c, self._ld(c), nb, row_first, row_step,
                                col_first, col_step)

    def getrf(self, a, ipiv, row0=0):              # in-place LU of the m x n view, pivots 1-based
        m, n = a.shape[1], a.shape[0]
        self.dev.getrf(m, n, a, self._ld(a), ipiv, self.info)
        self.first = torch.where((self.first == 0) & (self.info != 0), self.info + row0, self.first)

    def laswp(self, a, ipiv, k1, k2):              # rows k of a swapped with ipiv[k]-1, k1 <= k < k2
        m, n = a.shape[1], a.shape[0]
        self.dev.laswp(m, n, a, self._ld(a), k1, k2, ipiv)

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


def _reduce_info(info: int, device) -> int:
    """Smallest non-zero LAPACK info over the ranks."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return info
    big = 2 ** 31 - 1
    t = torch.tensor([info if info > 0 else big], dtype=torch.int64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    info = int(t.item())
    return 0 if info == big else info


def potrf_block_cyclic(a_local: torch.Tensor, n: int, nb: int, ops, grid: Grid | None = None):
    """In-place lower Cholesky of an n x n SPD matrix in 2-D block-cyclic layout.

    `a_local` has shape (nloc, mloc) — see `BlockCyclic`: the local column-major mloc x nloc
    matrix of this rank's row blocks x column blocks.  On return the blocks on or below the
    diagonal hold L (blocks above it, and the upper triangles of the diagonal blocks, are not
    meaningful).  Returns LAPACK's info (0, or the 1-based index of the first bad pivot).
    """
    grid = grid or Grid()
    lay = BlockCyclic(n, nb, grid)
    P, Q, p, q = grid.P, grid.Q, grid.p, grid.q
    nblk, mloc, nloc = lay.nblk, lay.mloc, lay.nloc
    assert tuple(a_local.shape) == (nloc, mloc), (tuple(a_local.shape), (nloc, mloc))
    dt, dv = a_local.dtype, a_local.device
    side = _SideStream(dv)
    # the two pieces of a panel this rank needs, double buffered by the parity of k
    row_bufs = [torch.empty(nb * max(mloc, 1), dtype=dt, device=dv) for _ in range(2)]
    col_bufs = [torch.empty(nb * max(nloc, 1), dtype=dt, device=dv) for _ in range(2)]
    diag_bufs = [torch.empty((nb, nb), dtype=dt, device=dv) for _ in range(2)]
    src_p = q % P            # the grid row whose panel rows contain this rank's column blocks

    def post(k):
        """Factor block column k (every earlier update applied) and start its exchange.
        Returns (lrow, lcol, event): lrow = panel rows of my row blocks > k as an (nb, mr) view,
        lcol = panel rows of my column blocks > k as an (nb, nc) view."""
        pk, qk = k % P, k % Q
        r0, c0 = lay.lrow_from(k + 1), lay.lcol_from(k + 1)
        mr, nc = mloc - r0, nloc - c0
        lrow = row_bufs[k % 2][: nb * mr].view(nb, mr)
        lcol = col_bufs[k % 2][: nb * nc].view(nb, nc)
        if k == nblk - 1:
            if p == pk and q == qk:
                lc, lr = (k - q) // Q * nb, (k - p) // P * nb
                ops.potrf_lower(a_local[lc: lc + nb, lr: lr + nb], k * nb)
            return lrow, lcol, None
        if q == qk:
            lc = (k - q) // Q * nb
            colblk = a_local[lc: lc + nb]                      # my rows of block column k
            d = diag_bufs[k % 2]
            if p == pk:
                lr = (k - p) // P * nb
                ops.potrf_lower(colblk[:, lr: lr + nb], k * nb)
                if P > 1 or mr > 0:
                    d.copy_(colblk[:, lr: lr + nb])
            if P > 1:
                dist.broadcast(d, src=grid.rank_of(pk, qk), group=grid.col_group)
            if mr > 0:
                ops.trsm_right_lower_ct(d, colblk[:, r0:])
                lrow.copy_(colblk[:, r0:])
        with side.section():
            if mr > 0 and Q > 1:
                dist.broadcast(lrow, src=grid.rank_of(p, qk), group=grid.row_group, async_op=True).wait()
            if nc > 0:
                if p == src_p:
                    # my column blocks j = q, q+Q, ... (> k) inside my row blocks p, p+P, ... (> k)
                    i_first = lay.row_blocks[r0 // nb]
                    j_first = lay.col_blocks[c0 // nb]
                    pos0, step = (j_first - i_first) // P, Q // P
                    lcol.view(nb, nc // nb, nb).copy_(lrow.view(nb, mr // nb, nb)[:, pos0::step, :])
                if P > 1:
                    dist.broadcast(lcol, src=grid.rank_of(src_p, q), group=grid.col_group,
                                   async_op=True).wait()
            ev = side.mark()
        return lrow, lcol, ev

    def update(k, lrow, lcol, g_lo, g_hi):
        """Apply panel k to my column blocks with global index in [g_lo, g_hi)."""
        r0, c0 = lay.lrow_from(k + 1), lay.lcol_from(k + 1)
        c_lo, c_hi = lay.lcol_from(g_lo), lay.lcol_from(g_hi)
        if c_hi <= c_lo:
            return
        gj = lay.col_blocks[c_lo // nb]                        # first column block of the range
        r1 = lay.lrow_from(gj)                                 # rows above it stay untouched
        if r1 >= mloc:
            return
        c = a_local[c_lo: c_hi][:, r1:]
        ops.gemm_nt_sub_stair(c, lrow[:, r1 - r0:], lcol[:, c_lo - c0: c_hi - c0], nb,
                              lay.row_blocks[r1 // nb], P, gj, Q)

    cur = post(0)
    for k in range(nblk - 1):
        lrow, lcol, ev = cur
        side.wait(ev)
        if q == (k + 1) % Q:
            update(k, lrow, lcol, k + 1, k + 2)                # look-ahead: next panel's column first
        cur = post(k + 1)
        update(k, lrow, lcol, k + 2, nblk)
    return _reduce_info(ops.info_value(), dv)


def getrf_block_cyclic(a_local: torch.Tensor, ipiv: torch.Tensor, n: int, nb: int, ops,
                       grid: Grid | None = None):
    """In-place LU with partial pivoting, P A = L U, of an n x n matrix whose block columns are
    dealt cyclically to the ranks (a 1 x Q grid: `a_local` has shape (nloc, n)).

    Every panel, and with it the pivot search of each column, stays on one GPU and is factored
    by the single-GPU recursive panel code, so the pivot sequence is the one LAPACK's dgetrf
    finds.  `ipiv` (int32, length n, on every rank) receives LAPACK's 1-based global pivot rows.
    Returns info (0, or the index of the first exactly zero pivot).
    """
    grid = grid or Grid(1, dist.get_world_size() if dist.is_initialized() else 1)
    if grid.P != 1:
        raise ValueError("getrf_block_cyclic distributes block columns only: use a 1 x Q grid")
    lay = BlockCyclic(n, nb, grid)
    Q, q, nblk, nloc = grid.Q, grid.q, lay.nblk, lay.nloc
    assert tuple(a_local.shape) == (nloc, n)
    dt, dv = a_local.dtype, a_local.device
    side = _SideStream(dv)
    bufs = [torch.empty(nb * n, dtype=dt, device=dv) for _ in range(2)]
    piv_bufs = [torch.zeros(nb, dtype=torch.int32, device=dv) for _ in range(2)]

    def post(k):
        """Factor panel k (all earlier updates applied) and start its broadcast."""
        k0 = k * nb
        panel = bufs[k % 2][: nb * (n - k0)].view(nb, n - k0)   # rows k0.. of block column k
        piv = piv_bufs[k % 2]
        if q == k % Q:
            lc = (k - q) // Q * nb
            blk = a_local[lc: lc + nb][:, k0:]
            ops.getrf(blk, piv, k0)
            panel.copy_(blk)
        ev = None
        if Q > 1:
            with side.section():
                w1 = dist.broadcast(panel, src=k % Q, group=grid.row_group, async_op=True)
                w2 = dist.broadcast(piv, src=k % Q, group=grid.row_group, async_op=True)
                w1.wait()
                w2.wait()
                ev = side.mark()
        return panel, piv, ev

    def apply(k, panel, piv, g_lo, g_hi, left=False):
        """Interchanges of panel k on my column blocks in [g_lo, g_hi); for blocks right of the
        panel also U12 = L11^-1 A12 and the rank-nb update of the rows below."""
        k0 = k * nb
        c_lo, c_hi = lay.lcol_from(g_lo), lay.lcol_from(g_hi)
        if c_hi <= c_lo:
            return
        cols = a_local[c_lo: c_hi]
        ops.laswp(cols[:, k0:], piv, 0, nb)                    # pivots are relative to row k0
        if left:
            return
        ops.trsm_left_lower_unit(panel[:, :nb], cols[:, k0: k0 + nb])
        if n - k0 - nb > 0:
            ops.gemm_nn_sub(cols[:, k0 + nb:], panel[:, nb:], cols[:, k0: k0 + nb])

    cur = post(0)
    for k in range(nblk):
        panel, piv, ev = cur
        side.wait(ev)
        ipiv[k * nb: (k + 1) * nb].copy_(piv + k * nb)
        if k + 1 < nblk:
            if q == (k + 1) % Q:
                apply(k, panel, piv, k + 1, k + 2)
            cur = post(k + 1)
        apply(k, panel, piv, k + 2, nblk)
        apply(k, panel, piv, 0, k, left=True)
    return _reduce_info(ops.info_value(), dv)


def _gather_blocks(parts, n, nb, nrhs, like):
    """x (nrhs, n) on every rank from the blocks {k: (nrhs, nb)} held by their owners."""
    x = torch.zeros((nrhs, n), dtype=like.dtype, device=like.device)
    for k, v in parts.items():
        x[:, k * nb: (k + 1) * nb] = v
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(x)
    return x


def potrs_block_cyclic(l_local: torch.Tensor, b: torch.Tensor, n: int, nb: int, ops,
                       grid: Grid | None = None):
    """Solve A X = B with the 2-D block-cyclic Cholesky factor left by `potrf_block_cyclic`
    (cho_solve, src/linalg/cholesky.cr:81-90 -> ?potrs).  `b` is the (nrhs, n) right-hand side,
    the same on every rank; returns X (nrhs, n) on every rank.

    Two block substitutions.  L Y = B: the contributions L_ik y_k are accumulated where L_ik lives;
    block k is solved by the owner of L_kk after a reduction along its grid row, and y_k goes down
    its grid column.  L^H X = Y: contributions L_jk^H x_j accumulate per column block, are reduced
    down the grid column, and x_k goes along the grid row.  Four small collectives per block: the
    solve is bound by their latency, not by bandwidth or flops.
    """
    grid = grid or Grid()
    lay = BlockCyclic(n, nb, grid)
    P, Q, p, q = grid.P, grid.Q, grid.p, grid.q
    nblk, mloc, nloc = lay.nblk, lay.mloc, lay.nloc
    nrhs = b.shape[0]
    acc = torch.zeros((nrhs, max(mloc, 1)), dtype=b.dtype, device=b.device)    # minus sum L_ik y_k
    y = {}
    for k in range(nblk):
        pk, qk = k % P, k % Q
        owner = grid.rank_of(pk, qk)
        t = None
        if p == pk:
            lr = (k - p) // P * nb
            t = acc[:, lr: lr + nb].clone()
            if Q > 1:
                dist.reduce(t, dst=owner, group=grid.row_group)
        if q == qk:
            lc = (k - q) // Q * nb
            yk = torch.empty((nrhs, nb), dtype=b.dtype, device=b.device)
            if p == pk:
                yk.copy_(b[:, k * nb: (k + 1) * nb] + t)
                ops.trsm_left(l_local[lc: lc + nb, lr: lr + nb], yk, lower=True)
                y[k] = yk
            if P > 1:
                dist.broadcast(yk, src=owner, group=grid.col_group)
            r0 = lay.lrow_from(k + 1)
            if mloc - r0 > 0:
                ops.gemm_nn_sub(acc[:, r0:], l_local[lc: lc + nb, r0:], yk)
    acc2 = torch.zeros((nrhs, max(nloc, 1)), dtype=b.dtype, device=b.device)   # minus sum L_jk^H x_j
    x = {}
    for k in range(nblk - 1, -1, -1):
        pk, qk = k % P, k % Q
        owner = grid.rank_of(pk, qk)
        t = None
        if q == qk:
            lc = (k - q) // Q * nb
            t = acc2[:, lc: lc + nb].clone()
            if P > 1:
                dist.reduce(t, dst=owner, group=grid.col_group)
        if p == pk:
            lr = (k - p) // P * nb
            xk = torch.empty((nrhs, nb), dtype=b.dtype, device=b.device)
            if q == qk:
                xk.copy_(y[k] + t)
                ops.trsm_left(l_local[lc: lc + nb, lr: lr + nb], xk, lower=True, conj_trans=True)
                x[k] = xk
            if Q > 1:
                dist.broadcast(xk, src=owner, group=grid.row_group)
            c1 = lay.lcol_from(k)                # my column blocks left of k
            if c1 > 0:
                ops.gemm_cn_sub(acc2[:, :c1], l_local[:c1, lr: lr + nb], xk)
    return _gather_blocks(x, n, nb, nrhs, b)


def getrs_block_cyclic(lu_local: torch.Tensor, ipiv: torch.Tensor, b: torch.Tensor, n: int, nb: int,
                       ops, grid: Grid | None = None):
    """Solve A X = B with the block-column LU left by `getrf_block_cyclic` (LUMatrix#solve,
    src/linalg/lu.cr:151-162 -> ?getrs 'N').  `b`: (nrhs, n), the same on every rank (it is
    overwritten by the row interchanges); returns X (nrhs, n) on every rank."""
    grid = grid or Grid(1, dist.get_world_size() if dist.is_initialized() else 1)
    if grid.P != 1:
        raise ValueError("getrs_block_cyclic works on block columns: use a 1 x Q grid")
    lay = BlockCyclic(n, nb, grid)
    Q, q, nblk = grid.Q, grid.q, lay.nblk
    nrhs = b.shape[0]
    ops.laswp(b, ipiv, 0, n)
    y, x = {}, {}
    for upper in (False, True):
        acc = torch.zeros((nrhs, n), dtype=b.dtype, device=b.device)       # minus the contributions
        order = range(nblk - 1, -1, -1) if upper else range(nblk)
        for k in order:
            k0 = k * nb
            t = acc[:, k0: k0 + nb].clone()
            if Q > 1:
                dist.reduce(t, dst=k % Q, group=grid.row_group)
            if q == k % Q:
                lc = (k - q) // Q * nb
                blk = lu_local[lc: lc + nb]                                    # my block column k
                v = ((y[k] if upper else b[:, k0: k0 + nb]) + t).contiguous()
                ops.trsm_left(blk[:, k0: k0 + nb], v, lower=not upper, unit=not upper)
                (x if upper else y)[k] = v
                if upper and k0 > 0:
                    ops.gemm_nn_sub(acc[:, :k0], blk[:, :k0], v)
                if not upper and n - k0 - nb > 0:
                    ops.gemm_nn_sub(acc[:, k0 + nb:], blk[:, k0 + nb:], v)
    return _gather_blocks(x, n, nb, nrhs, b)


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


class HostGemmPlan:
    """C <- A B + C for a product split by block columns over the ranks, operands in (pinned)
    HOST memory on every rank: rank r holds its block columns of B and C and only the r-th slice
    of A's columns.  Each rank uploads its slice of A over its own PCIe link and the slices are
    all-gathered over NVLink, instead of every rank pulling the whole of A through the host
    (world x the PCIe traffic of the replicated layout).  Device buffers are allocated once per
    plan and reused by every `run`."""

    def __init__(self, n: int, nc: int, dtype=torch.float64, group=None):
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.group, self.n, self.nc = group, n, nc
        if n % self.world != 0:
            raise ValueError("n must be a multiple of the number of ranks")
        self.a_full = torch.empty((n, n), dtype=dtype, device="cuda")
        self.b = torch.empty((nc, n), dtype=dtype, device="cuda")
        self.c = torch.empty((nc, n), dtype=dtype, device="cuda")
        self.copy = torch.cuda.Stream()
        self.down = torch.cuda.Stream()
        from . import dev
        self.dev = dev

    def a_slice(self):
        """Column range [c0, c1) of A this rank must supply."""
        w = self.n // self.world
        return self.rank * w, (self.rank + 1) * w

    def run(self, hA_slice: torch.Tensor, hB: torch.Tensor, hC: torch.Tensor, nchunks: int = 8):
        """hA_slice: (n/world, n) host tensor = this rank's columns of A; hB, hC: (nc, n) host
        tensors.  On return (host-synchronous) hC holds the result.

        A's slice goes up first and is gathered over NVLink while B and C follow in column
        chunks on the copy stream; chunk j is multiplied as soon as it has landed and travels
        home on a third stream while chunk j+1 is multiplied."""
        n, nc, dev = self.n, self.nc, self.dev
        c0, c1 = self.a_slice()
        cur = torch.cuda.current_stream()
        self.copy.wait_stream(cur)
        self.down.wait_stream(cur)
        self.a_full[c0:c1].copy_(hA_slice, non_blocking=True)
        work = None
        if self.world > 1:
            work = dist.all_gather_into_tensor(self.a_full, self.a_full[c0:c1], group=self.group,
                                               async_op=True)
        step = max(128, -(-nc // nchunks) // 128 * 128)
        chunks = [(j, min(nc, j + step)) for j in range(0, nc, step)]
        ups = []
        with torch.cuda.stream(self.copy):
            for j0, j1 in chunks:
                self.b[j0:j1].copy_(hB[j0:j1], non_blocking=True)
                self.c[j0:j1].copy_(hC[j0:j1], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(self.copy)
                ups.append(ev)
        if work is not None:
            work.wait()
        for (j0, j1), ev in zip(chunks, ups):
            cur.wait_event(ev)
            dev.gemm(dev.NO_TRANS, dev.NO_TRANS, n, j1 - j0, n, 1.0, self.a_full, n, self.b[j0:j1], n,
                     1.0, self.c[j0:j1], n)
            done = torch.cuda.Event()
            done.record(cur)
            with torch.cuda.stream(self.down):
                self.down.wait_event(done)
                hC[j0:j1].copy_(self.c[j0:j1], non_blocking=True)
        cur.wait_stream(self.down)
        cur.synchronize()
