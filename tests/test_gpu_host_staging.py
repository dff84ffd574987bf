"""GPU tests of the pageable-host transfer path of the drop-in symbols (host_stage.cu).

The reference hands the library slices of its GC heap (src/matrix/general_matrix.cr:26), i.e.
pageable memory; transfers of 4 MiB and more then go through the library's ring of pinned slots.
The results must be the same bits as with pinned buffers (which take the direct DMA path), for
padded leading dimensions, for columns longer than one slot and from several threads at once."""
import threading

import numpy as np
import pytest
import torch

from linalg_b200 import ffi
from oracle import checks

pytestmark = pytest.mark.gpu


def _pinned_like(a):
    """A pinned copy of a Fortran-ordered numpy array (same strides)."""
    t = torch.empty(a.T.shape, dtype=torch.from_numpy(np.zeros(1, a.dtype)).dtype).pin_memory()
    v = t.numpy().T
    v[...] = a
    return t, v


def _dgemm(abi, A, B, C, m, n, k, lda, ldb, ldc):
    abi.gemm("d")(ffi.COL_MAJOR, ffi.NO_TRANS, ffi.NO_TRANS, m, n, k, 1.0, ffi.ptr(A), lda, ffi.ptr(B), ldb,
                  1.0, ffi.ptr(C), ldc)


def test_pageable_and_pinned_operands_give_identical_results(product_abi, oracle_abi):
    rng = np.random.default_rng(11)
    m, n, k = 3000, 2600, 1100
    lda, ldb, ldc = m + 5, k + 3, m + 8           # padded leading dimensions
    A = np.asfortranarray(rng.random((lda, k)) - 0.5)
    B = np.asfortranarray(rng.random((ldb, n)) - 0.5)
    C0 = np.asfortranarray(rng.random((ldc, n)) - 0.5)
    Cp = C0.copy(order="F")
    _dgemm(product_abi, A, B, Cp, m, n, k, lda, ldb, ldc)        # pageable: staged
    keep = [_pinned_like(x) for x in (A, B, C0)]
    (_, Ah), (_, Bh), (_, Ch) = keep
    _dgemm(product_abi, Ah, Bh, Ch, m, n, k, lda, ldb, ldc)      # pinned: direct
    assert np.array_equal(Cp, Ch)
    assert np.array_equal(Cp[m:], C0[m:])                          # padding rows untouched
    ref = C0.copy(order="F")
    _dgemm(oracle_abi, A, B, ref, m, n, k, lda, ldb, ldc)
    assert checks.rel(Cp[:m] - ref[:m], ref[:m]) < 1e-13


def test_columns_longer_than_a_staging_slot(product_abi):
    """A 2^22-row column is 32 MiB, two slots: tall-skinny least squares from pageable memory."""
    rng = np.random.default_rng(12)
    m, n = 1 << 22, 6
    A = np.asfortranarray(rng.random((m, n)))
    x = rng.random((n, 1))
    b = np.asfortranarray(A @ x)
    from linalg_b200 import la
    with la.use_backend(product_abi):
        got = la.GeneralMatrix(A).solvels(la.GeneralMatrix(b))
    assert np.abs(got.raw[:n] - x).max() < 1e-9


def test_factorisations_from_pageable_memory_match_the_oracle(product_abi, oracle_abi):
    from linalg_b200 import la
    rng = np.random.default_rng(13)
    n = 4200                                                       # streams blocks home while factoring
    A = np.asfortranarray(rng.random((n, n)))
    G = rng.random((n, n)) - 0.5
    S = np.asfortranarray(G @ G.T / n + np.eye(n))
    out = {}
    for name, abi in (("gpu", product_abi), ("cpu", oracle_abi)):
        with la.use_backend(abi):
            lu = la.GeneralMatrix(A).lu_factor()
            ch = la.GeneralMatrix(S).cholesky(lower=True)
            out[name] = (lu.a.raw.copy(), np.asarray(lu.ipiv).copy(), ch.raw.copy())
    assert np.array_equal(out["gpu"][1], out["cpu"][1])
    assert checks.lu_residual(A, out["gpu"][0], out["gpu"][1]) < 1e-12
    assert checks.chol_residual(S, out["gpu"][2], "L") < 1e-12
    assert checks.rel(np.tril(out["gpu"][2]) - np.tril(out["cpu"][2]), out["cpu"][2]) < 1e-12


def test_concurrent_callers_with_staged_transfers(product_abi, oracle_abi):
    rng = np.random.default_rng(14)
    n = 1500
    jobs = [(np.asfortranarray(rng.random((n, n))), np.asfortranarray(rng.random((n, n)))) for _ in range(6)]
    got = [np.zeros((n, n), order="F") for _ in jobs]
    errs = []

    def work(i):
        try:
            _dgemm(product_abi, jobs[i][0], jobs[i][1], got[i], n, n, n, n, n, n)
        except Exception as e:  # noqa: BLE001
            errs.append(e)

    ts = [threading.Thread(target=work, args=(i,)) for i in range(len(jobs))]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errs, errs
    for (A, B), C in zip(jobs, got):
        ref = np.zeros((n, n), order="F")
        _dgemm(oracle_abi, A, B, ref, n, n, n, n, n, n)
        assert checks.rel(C - ref, ref) < 1e-13


def test_drop_in_dgemm_split_over_the_visible_gpus():
    """LA_B200_GPUS > 1: the columns of a big cblas_dgemm are dealt to one host thread per GPU
    (shim_cblas.cu).  The switch is read when the library is loaded, so the call runs in a child
    process; with one visible GPU the request is clamped and the ordinary path is checked."""
    import os
    import subprocess
    import sys
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent
    env = dict(os.environ, LA_B200_GPUS="2")
    r = subprocess.run([sys.executable, str(root / "tools" / "multi_gpu_gemm_check.py"), "4096"],
                       cwd=str(root), env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "max relative error on sampled columns" in r.stdout
