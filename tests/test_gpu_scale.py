"""GPU tests at the sizes BASELINE.json quotes, through size-independent properties.

The small parity cases cannot see scheduling-dependent faults (a tile handed back to the TMA
producer while a slow warp is still reading it shows up only with thousands of tiles in flight),
so these tests run the device-resident entry points at scale and check
  * bit-for-bit repeatability of the product over many launches (any race breaks it),
  * the product against an independent FP64 product (torch, used as a checker only),
  * ||P A X - L U X|| and ||A X - L L^H X|| for random probe blocks X.
"""
import numpy as np
import pytest
import torch

from linalg_b200 import dev

pytestmark = pytest.mark.gpu


def _rand(shape, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return torch.rand(shape, dtype=torch.float64, device="cuda", generator=g) - 0.5


@pytest.mark.parametrize("ta,tb,m,n,k", [("N", "N", 8000, 7000, 512), ("N", "T", 6000, 6000, 1024),
                                         ("T", "N", 4096, 5120, 768)])
def test_dgemm_is_repeatable_and_right_at_scale(ta, tb, m, n, k):
    op = {"N": dev.NO_TRANS, "T": dev.TRANS}
    ar, ac = (m, k) if ta == "N" else (k, m)
    br, bc = (k, n) if tb == "N" else (n, k)
    A, B, C0 = _rand((ac, ar), 1), _rand((bc, br), 2), _rand((n, m), 3)
    Am = A.T if ta == "N" else A
    Bm = B.T if tb == "N" else B
    ref = (C0.T - Am @ Bm).T
    first = None
    for rep in range(12):
        C = C0.clone()
        dev.gemm(op[ta], op[tb], m, n, k, -1.0, A, ar, B, br, 1.0, C, m)
        if first is None:
            first = C
            assert dev.last_gemm_path() == 1          # the FP64 tensor-core kernel
            assert (C - ref).abs().max().item() < 1e-13 * k
        else:
            assert torch.equal(C, first), f"launch {rep} differs from launch 0"


@pytest.mark.parametrize("ta,tb,m,n,k,alpha,beta", [
    ("N", "N", 1792, 1792, 1792, 1.0, 0.0),      # 196 tiles on 148 SMs: the last 48 are split along k
    ("T", "N", 2304, 2304, 2304, -0.5, 1.0),
    ("N", "T", 3840, 3700, 1500, 2.0, -1.5),     # ragged edge tiles among the split ones
    ("N", "N", 200, 300, 40000, 1.0, 0.25)])     # few tiles, long k: every tile split
def test_dgemm_split_tiles_are_repeatable_and_right(ta, tb, m, n, k, alpha, beta):
    # products whose tiles do not fill the SMs a whole number of times cut the tiles of the last
    # wave (or all of them) along k and reduce the pieces in a fixed order
    op = {"N": dev.NO_TRANS, "T": dev.TRANS}
    ar, ac = (m, k) if ta == "N" else (k, m)
    br, bc = (k, n) if tb == "N" else (n, k)
    A, B, C0 = _rand((ac, ar), 31), _rand((bc, br), 32), _rand((n, m), 33)
    Am = A.T if ta == "N" else A
    Bm = B.T if tb == "N" else B
    ref = (beta * C0.T + alpha * (Am @ Bm)).T
    first = None
    for rep in range(6):
        C = C0.clone()
        dev.gemm(op[ta], op[tb], m, n, k, alpha, A, ar, B, br, beta, C, m)
        if first is None:
            first = C
            assert dev.last_gemm_path() == 1
            assert (C - ref).abs().max().item() < 2e-14 * k
        else:
            assert torch.equal(C, first), f"launch {rep} differs from launch 0"


def test_masked_rank_k_update_is_repeatable_and_right_at_scale():
    n, k = 15360, 512
    A, C0 = _rand((k, n), 4), _rand((n, n), 5)
    low = torch.tril(torch.ones((n, n), dtype=torch.bool, device="cuda")).T    # torch view of "lower"
    first = None
    for rep in range(6):
        C = C0.clone()
        dev.herk(dev.LOWER, dev.NO_TRANS, n, k, -1.0, A, n, 1.0, C, n)
        if first is None:
            first = C
            ref = (C0.T - A.T @ A).T
            assert ((C - ref).abs() * low).max().item() < 1e-13 * k
            assert torch.equal(C[~low], C0[~low])      # the other triangle is never touched
            del ref
        else:
            assert torch.equal(C, first), f"launch {rep} differs from launch 0"


@pytest.mark.parametrize("n", [8192, 16384])
def test_getrf_probe_residual_at_scale(n):
    A0 = _rand((n, n), 6) + 0.5                       # torch (cols, rows) view of column-major A
    W = A0.clone()
    ipiv = torch.zeros(n, dtype=torch.int32, device="cuda")
    info = torch.zeros(1, dtype=torch.int32, device="cuda")
    dev.getrf(n, n, W, n, ipiv, info)
    assert info.item() == 0
    piv = ipiv.cpu().numpy().astype(np.int64) - 1
    assert piv.min() >= 0 and piv.max() < n and np.all(piv >= np.arange(n))
    perm = np.arange(n)
    for i, p in enumerate(piv):
        perm[i], perm[p] = perm[p], perm[i]
    X = _rand((n, 16), 7)                              # math (n, 16)
    M = W.T                                            # math view of the factors
    LUx = torch.tril(M, -1) @ (torch.triu(M) @ X) + torch.triu(M) @ X
    PAx = (A0.T @ X)[torch.from_numpy(perm).cuda()]
    rel = ((PAx - LUx).abs().max() / (A0.abs().max() * X.abs().sum(dim=0).max())).item()
    assert rel < 1e-12, rel
    assert torch.tril(M, -1).abs().max().item() <= 1.0 + 1e-14     # partial pivoting: |l_ij| <= 1


@pytest.mark.parametrize("m,n", [(8192, 8192), (16384, 16384), (12000, 9000), (9000, 12000)])
def test_sgetrf_probe_residual_at_scale(m, n):
    # FP32 LU with look-ahead (the 3xTF32 trailing update beside the next panel on a side stream)
    g = torch.Generator(device="cuda").manual_seed(60 + m % 7)
    A0 = torch.rand((n, m), dtype=torch.float32, device="cuda", generator=g)     # torch view of m x n
    W = A0.clone()
    k = min(m, n)
    ipiv = torch.zeros(k, dtype=torch.int32, device="cuda")
    info = torch.zeros(1, dtype=torch.int32, device="cuda")
    dev.getrf(m, n, W, m, ipiv, info)
    assert info.item() == 0
    piv = ipiv.cpu().numpy().astype(np.int64) - 1
    assert piv.min() >= 0 and piv.max() < m and np.all(piv >= np.arange(k))
    perm = np.arange(m)
    for i, p in enumerate(piv):
        perm[i], perm[p] = perm[p], perm[i]
    X = torch.rand((n, 8), dtype=torch.float64, device="cuda", generator=g) - 0.5
    M = W.T.double()                                   # math view of the factors (m x n)
    L = torch.tril(M[:, :k], -1) + torch.eye(m, k, dtype=torch.float64, device="cuda")
    U = torch.triu(M[:k, :])
    PAx = (A0.T.double() @ X)[torch.from_numpy(perm).cuda()]
    rel = ((PAx - L @ (U @ X)).abs().max() / (A0.abs().max().double() * X.abs().sum(dim=0).max())).item()
    assert rel < 2e-5, rel
    assert torch.tril(M[:, :k], -1).abs().max().item() <= 1.0 + 1e-6


@pytest.mark.parametrize("n", [8192, 16384])
def test_potrf_probe_residual_at_scale(n):
    G = _rand((n, n), 8)
    S = (G + G.T) / 2 + n * torch.eye(n, dtype=torch.float64, device="cuda")
    del G
    W = S.clone()
    info = torch.zeros(1, dtype=torch.int32, device="cuda")
    dev.potrf(dev.LOWER, n, W, n, info)
    assert info.item() == 0
    assert torch.equal(torch.triu(W.T, 1), torch.triu(S.T, 1))      # strictly upper part untouched
    L = torch.tril(W.T)
    X = _rand((n, 16), 9)
    rel = ((S @ X - L @ (L.T @ X)).abs().max() / (S.abs().max() * X.abs().sum(dim=0).max())).item()
    assert rel < 1e-13, rel


@pytest.mark.parametrize("n,lower", [(4608, True), (4700, False), (8300, True), (8300, False)])
def test_host_potrf_moves_only_the_named_triangle(n, lower, product_abi):
    # the host-pointer dpotrf_ uploads and downloads only the named triangle (in block columns
    # that match the factorisation's blocks) and streams finished blocks home while it works:
    # the other triangle of the caller's array must come back bit-identical, poison included
    import ctypes as C
    from linalg_b200 import ffi
    rng = np.random.default_rng(12)
    G = rng.random((n, n)) - 0.5
    S = G @ G.T / n + np.eye(n)
    A = np.asfortranarray(S.copy())
    other = np.triu_indices(n, 1) if lower else np.tril_indices(n, -1)
    A[other] = np.nan
    before = A.copy(order="F")
    info = C.c_int(-99)
    ni = C.c_int(n)
    product_abi.lapack("potrf", "d")(C.c_char_p(b"L" if lower else b"U"), C.byref(ni), ffi.ptr(A),
                                     C.byref(ni), C.byref(info))
    assert info.value == 0
    assert np.array_equal(np.isnan(A), np.isnan(before))          # poison where it was, only there
    F = np.tril(A) if lower else np.triu(np.nan_to_num(A))
    F = np.nan_to_num(F)
    R = F @ F.T if lower else F.T @ F
    assert np.abs(R - S).max() / np.abs(S).max() < 1e-13


@pytest.mark.parametrize("m,n", [(4608, 4608), (8300, 8300), (9000, 5000), (16384, 16384)])
def test_host_getrf_streams_block_rows_home(m, n, product_abi, oracle_abi):
    # the host-pointer dgetrf_ sends every finished block row home while it is still factoring
    # (shim_lapack.cu); factors and pivots must equal LAPACK's as for the small cases.  16384 is
    # the benchmark size: look-ahead on the side stream, standard and wide leaves, split steps
    import ctypes as C
    from linalg_b200 import ffi
    rng = np.random.default_rng(13)
    A0 = np.asfortranarray(rng.random((m, n)))
    out = []
    for abi in (product_abi, oracle_abi):
        A = A0.copy(order="F")
        ipiv = np.zeros(min(m, n), dtype=np.int32)
        info = C.c_int(-99)
        mi, ni = C.c_int(m), C.c_int(n)
        abi.lapack("getrf", "d")(C.byref(mi), C.byref(ni), ffi.ptr(A), C.byref(mi), ffi.ptr(ipiv),
                                 C.byref(info))
        assert info.value == 0
        out.append((A, ipiv))
    (lu, piv), (rlu, rpiv) = out
    assert np.array_equal(piv, rpiv)
    assert np.abs(lu - rlu).max() / np.abs(rlu).max() < 1e-10


@pytest.mark.parametrize("side,uplo,op,diag,na", [
    ("R", "L", "C", "N", 128), ("L", "L", "N", "U", 128), ("L", "U", "N", "N", 100),
    ("R", "U", "N", "N", 77), ("L", "L", "T", "N", 128), ("R", "L", "N", "U", 96)])
def test_trsm_wide_tensor_core_base_case(side, uplo, op, diag, na):
    # triangles of 65..128 with thousands of vectors take the 128-wide DMMA base kernel
    nvec = 6000
    g = torch.Generator(device="cuda").manual_seed(21)
    T = torch.rand((na, na), dtype=torch.float64, device="cuda", generator=g) - 0.5
    T += na * torch.eye(na, dtype=torch.float64, device="cuda")
    Tm = T.T                                             # math view of the column-major triangle
    tri = torch.tril(Tm) if uplo == "L" else torch.triu(Tm)
    if diag == "U":
        tri = tri - torch.diag(torch.diag(tri)) + torch.eye(na, dtype=torch.float64, device="cuda")
    M = tri if op == "N" else tri.T
    if side == "L":
        B = torch.rand((nvec, na), dtype=torch.float64, device="cuda", generator=g)   # na x nvec
        X = B.clone()
        dev.trsm(dev.LEFT, dev.LOWER if uplo == "L" else dev.UPPER,
                 {"N": dev.NO_TRANS, "T": dev.TRANS, "C": dev.CONJ_TRANS}[op],
                 dev.UNIT if diag == "U" else dev.NON_UNIT, na, nvec, 2.0, T, na, X, na)
        res = M @ X.T - 2.0 * B.T
    else:
        B = torch.rand((na, nvec), dtype=torch.float64, device="cuda", generator=g)   # nvec x na
        X = B.clone()
        dev.trsm(dev.RIGHT, dev.LOWER if uplo == "L" else dev.UPPER,
                 {"N": dev.NO_TRANS, "T": dev.TRANS, "C": dev.CONJ_TRANS}[op],
                 dev.UNIT if diag == "U" else dev.NON_UNIT, nvec, na, 2.0, T, na, X, nvec)
        res = X.T @ M - 2.0 * B.T
    assert res.abs().max().item() < 1e-11


def test_device_api_calls_on_two_streams_do_not_share_scratch_in_flight():
    """One host thread, two streams: every device-API call reuses the thread's scratch arena
    (hi/lo operand packs of the 3xTF32 product, LU exchange area), so a call on another stream
    must be ordered behind the previous one.  Alternating streams without any host-side
    synchronisation has to give the results of the serial run, bit for bit."""
    import torch
    from linalg_b200 import dev, ffi
    n = 2048
    g = torch.Generator(device="cuda").manual_seed(5)
    mats = [torch.rand((n, n), dtype=torch.float32, device="cuda", generator=g) for _ in range(4)]
    lus = [torch.rand((n, n), dtype=torch.float64, device="cuda", generator=g) for _ in range(2)]

    def run(streams):
        outs = [torch.zeros((n, n), dtype=torch.float32, device="cuda") for _ in range(4)]
        facs = [x.clone() for x in lus]
        piv = [torch.zeros(n, dtype=torch.int32, device="cuda") for _ in range(2)]
        info = torch.zeros(1, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        for rep in range(3):
            for i in range(4):
                with torch.cuda.stream(streams[i % len(streams)]):
                    dev.gemm(ffi.NO_TRANS, ffi.TRANS, n, n, n, 1.0, mats[i], n, mats[(i + 1) % 4], n,
                             0.0, outs[i], n)
        for i in range(2):
            with torch.cuda.stream(streams[i % len(streams)]):
                dev.getrf(n, n, facs[i], n, piv[i], info)
        torch.cuda.synchronize()
        return outs + facs + piv

    serial = run([torch.cuda.current_stream()])
    mixed = run([torch.cuda.Stream(), torch.cuda.Stream()])
    for a, b in zip(serial, mixed):
        assert torch.equal(a, b)


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32, torch.complex128])
@pytest.mark.parametrize("p,q,m", [(8, 8, 100000), (16, 5, 65537), (32, 32, 200001), (64, 64, 70000),
                                   (33, 64, 50000), (7, 24, 300000)])
def test_tall_gram_products_match_a_double_precision_reference(dtype, p, q, m):
    """X^H Y with a tiny result and a long reduction (the block-reflector products of the
    tall-skinny QR) takes the streaming kernel: compare with torch in double precision, and make
    sure two runs give the same bits (the partial sums are added in a fixed order)."""
    from linalg_b200 import dev, ffi
    g = torch.Generator(device="cuda").manual_seed(17)
    ldx, ldy = (m + 7) // 8 * 8 + 8, (m + 7) // 8 * 8   # 16-byte aligned columns for every type
    X = torch.rand((p, ldx), dtype=dtype, device="cuda", generator=g) - 0.5
    Y = torch.rand((q, ldy), dtype=dtype, device="cuda", generator=g) - 0.5
    S = torch.full((q, p), 7.0, dtype=dtype, device="cuda")
    S2 = torch.full((q, p), -3.0, dtype=dtype, device="cuda")
    dev.gemm(ffi.CONJ_TRANS, ffi.NO_TRANS, p, q, m, 1.0, X, ldx, Y, ldy, 0.0, S, p)
    path = dev.last_gemm_path()
    dev.gemm(ffi.CONJ_TRANS, ffi.NO_TRANS, p, q, m, 1.0, X, ldx, Y, ldy, 0.0, S2, p)
    assert torch.equal(S, S2)
    hi = torch.complex128 if dtype.is_complex else torch.float64
    ref = (X[:, :m].to(hi).conj() @ Y[:, :m].to(hi).T).T        # (q, p) = column-major p x q
    tol = 2e-4 if dtype == torch.float32 else 1e-12
    assert float((S.to(hi) - ref).abs().max()) <= tol * float(ref.abs().max())
    if not (dtype.is_complex and max(p, q) > 32):
        assert path == 4, "the tall-gram kernel was not taken"


@pytest.mark.parametrize("dt", [torch.float64, torch.float32, torch.complex128], ids=["d", "s", "z"])
@pytest.mark.parametrize("m,n,pad", [(300, 64, 8), (1500, 96, 3), (9000, 40, 8), (20000, 33, 5), (4100, 700, 16)])
def test_getrf_leaves_padding_rows_alone(dt, m, n, pad):
    # the LU leaf writes every row to its final position once, at the end of the launch (rows do
    # not move in between): with lda > m the rows between m and lda must come back untouched, and
    # the factors must reproduce P A (all leaf forms: one CTA, 512-thread and 256-thread CTAs,
    # several rows per thread)
    lda = m + pad
    g = torch.Generator(device="cuda").manual_seed(m + n)
    full = torch.rand((n, lda), dtype=dt, device="cuda", generator=g) - 0.5     # torch view: n x lda
    full[:, m:] = float("nan")
    A0 = full[:, :m].clone()
    W = full.clone()
    k = min(m, n)
    ipiv = torch.zeros(k, dtype=torch.int32, device="cuda")
    info = torch.zeros(1, dtype=torch.int32, device="cuda")
    dev.getrf(m, n, W, lda, ipiv, info)
    assert info.item() == 0
    assert torch.isnan(W[:, m:].real).all() and not torch.isnan(W[:, :m].real).any()
    piv = ipiv.cpu().numpy().astype(np.int64) - 1
    perm = np.arange(m)
    for i, p in enumerate(piv):
        perm[i], perm[p] = perm[p], perm[i]
    hp = torch.complex128 if dt == torch.complex128 else torch.float64
    M = W[:, :m].T.to(hp)                                                        # m x n factors
    L = torch.tril(M[:, :k], -1) + torch.eye(m, k, dtype=hp, device="cuda")
    U = torch.triu(M[:k, :])
    PA = A0.T.to(hp)[torch.from_numpy(perm).cuda()]
    tol = 5e-5 if dt == torch.float32 else 1e-12
    assert ((PA - L @ U).abs().max() / A0.abs().max()).item() < tol * max(1, n // 50)
