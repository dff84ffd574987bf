#!/usr/bin/env python
"""bench.py — headline benchmark of the la_b200 backend (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--n 16384]

Metric: FP64 GEMM TFLOP/s at N = 16384 (the DGEMM leg of "DGEMM/DGETRF/DPOTRF TFLOP/s at
N=16384"), in the reference's `Matrix#*` call pattern: C <- 1.0*A*B + 1.0*C on a zeroed C,
column-major, NoTrans/NoTrans (src/linalg/mult.cr:26,64-69,119-120).  One step = one such
product.  DGETRF / DPOTRF at the same N are timed in the same run and reported under "extra",
each with its own roofline object.

  value   device-resident: operands already in HBM, CUDA events on the launching stream
          (la_b200_dgemm_dev), max over ranks.
  e2e     N = 1: the drop-in symbol cblas_dgemm called with HOST pointers (pinned), i.e. H2D of
          A, B, C + kernel + D2H of C inside the timed region; "e2e_pageable" is the same call on
          plain (unpinned) numpy buffers, which is what the reference's GC-heap slices are
          (src/matrix/general_matrix.cr:26).  N > 1: `dist.HostGemmPlan` — every rank uploads its
          block columns of B and C and 1/N of A from pinned host memory, A is all-gathered over
          NVLink, the result block goes back to the host.
  N > 1   strong scaling of the product: B and C are split by block columns over the ranks
          (SURVEY.md §8e), no collective on the device-resident data path.  The factorisations
          of BASELINE config c4 / the scaling target run in the same job: Cholesky N = 65536 on
          the 2-D block-cyclic process grid and LU N = 65536 on block columns
          (linalg_b200/dist.py), each with its distributed residual and its speed-up over the
          single-GPU routine timed on rank 0 in the same run, plus the TSQR least squares
          (2^20 x 512) — all under "extra.dist_*".
  --impl reference   the reference's own arithmetic path — OpenBLAS dgemm through the same C ABI
          on the host cores (all threads) — at the SAME N.

Inputs are synthetic U[0,1) (Mat.rand, src/matrix/matrix.cr:490-492).  Operands (3 x 2 GiB)
exceed the 126 MB L2, so no explicit flush is needed between iterations.
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

FP64_DMMA_PEAK_TFLOPS = 37.0   # measured: tools/microbench/peak_fp64.cu on this pool's B200
# measured: tools/microbench/peak_tf32.cu (tcgen05.mma.kind::tf32, M128 N256 K8, TMEM resident),
# profiles/r2_peak_tf32.jsonl: best single launch / back to back for 3 s
TF32_PEAK_BURST_TFLOPS = 1115.1
TF32_PEAK_SUSTAINED_TFLOPS = 1018.7
DGEMM_KCHUNK = 4096                     # the library runs long-k products as k-chunks (gemm_dmma.cu)
DGEMM_DRAM_BYTES_PER_LAUNCH = 18.82e9   # one k-chunk launch at N = 16384: 16.69 GB read + 2.14 GB written (ncu)
WORKLOAD = "dgemm N={n} NoTrans/NoTrans alpha=1 beta=1 on zeroed C (Matrix#*)"
DIST_N = 65536                          # BASELINE config c4 / north_star scaling target
DIST_NB = 512
DIST_RESERVE_SMS = int(os.environ.get("LA_DIST_RESERVE_SMS", "8"))
DIST_COMM_CTAS = int(os.environ.get("LA_DIST_COMM_CTAS", "8"))
# kernel time shares of one factorisation at N = 16384 from the ncu launch lists (serialised,
# cold caches: shares, not absolutes); refreshed whenever the kernels change
FACTOR_PROFILE = {
    "dgetrf": {"source": "profiles/r2g_dgetrf16384_launches.csv",
               "kernels_ms": {"dgemm_dmma_kernel (trailing + in-panel updates)": 119.9,
                              "lu_panel_kernel (32- and 64-column leaves)": 35.3, "laswp_*": 13.3,
                              "trsm_dmma_kernel": 12.3}},
    "dpotrf": {"source": "profiles/r2g_dpotrf16384_launches.csv",
               "kernels_ms": {"dgemm_dmma_kernel (HERK / block-column updates)": 53.8,
                              "potf2_kernel (128-column leaves)": 16.9, "trsm_dmma_kernel": 6.3}},
}


def _peaks():
    try:
        return json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
    except Exception:
        return {}


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""

    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index=0):
        self.index = index
        self.rows = []
        self._stop = threading.Event()
        self._t = None

    def _run(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}",
                                      "--format=csv,noheader,nounits"], capture_output=True,
                                     text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.2)

    def __enter__(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self):
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for nm, v in zip(self.NAMES, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                continue
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------- CPU arm
def _cpu_gemm_setup(n):
    from linalg_b200 import ffi
    from oracle import openblas_ref
    cores = len(os.sched_getaffinity(0))
    openblas_ref.set_num_threads(cores)
    rng = np.random.default_rng(20261106)
    A = np.asfortranarray(rng.random((n, n)))
    B = np.asfortranarray(rng.random((n, n)))
    Cm = np.zeros((n, n), order="F")
    f = openblas_ref.abi().gemm("d")

    def step():
        Cm[:] = 0
        f(ffi.COL_MAJOR, ffi.NO_TRANS, ffi.NO_TRANS, n, n, n, 1.0, ffi.ptr(A), n, ffi.ptr(B), n, 1.0,
          ffi.ptr(Cm), n)

    return step, cores, openblas_ref.config()


def cpu_dgemm_tflops(n, reps=2):
    """OpenBLAS dgemm (the library the reference links) on all host cores; best of reps."""
    step, cores, cfg = _cpu_gemm_setup(n)
    best = 1e30
    for _ in range(reps + 1):
        t0 = time.perf_counter()
        step()
        best = min(best, time.perf_counter() - t0)
    return 2.0 * n ** 3 / best * 1e-12, cores, cfg, reps + 1


def run_reference(args):
    """Reference arm: the reference's CPU arithmetic path on the box's host cores (rank 0), on
    the same workload as the GPU arm (dgemm at N = args.n, ~6 s per step on 16 cores)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    n = args.n
    step, cores, cfg = _cpu_gemm_setup(n)
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = (time.perf_counter() - t0) / args.steps
    tf = 2.0 * n ** 3 / dt * 1e-12
    sample = f"the full workload: dgemm N={n} (2N^3 = {2.0 * n ** 3:.3e} flop) per step"
    print(json.dumps({
        "impl": "reference", "metric": "fp64_gemm_tflops", "value": tf, "unit": "TFLOP/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": WORKLOAD.format(n=n)},
        "cpu_baseline": {"value": tf, "unit": "TFLOP/s", "cores": cores, "kind": "reference",
                         "sample": sample + "; " + cfg},
        "e2e": {"value": tf, "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ------------------------------------------------------------------------- GPU legs
def _best_ms(fn, reps=3, setup=None):
    import torch
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    b = 1e30
    for i in range(reps + 1):
        if setup:
            setup()
        torch.cuda.synchronize()
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        if i:
            b = min(b, e0.elapsed_time(e1))
    return b


def _roofline_obj(name, tflops, ms, peak):
    prof = FACTOR_PROFILE[name]
    tot = sum(prof["kernels_ms"].values())
    return {"bound": "tensor", "achieved": tflops, "peak": peak, "unit": "TFLOP/s",
            "frac": tflops / peak, "ms": ms,
            "kernel_share": {k: round(v / tot, 3) for k, v in prof["kernels_ms"].items()},
            "kernel_share_source": prof["source"] + " (ncu launch list, serialised: shares only)"}


def time_factorizations(n, steps):
    """DGETRF and DPOTRF at the same N, device resident, with probe-vector parity checks, and the
    s/z factorisations at the same N."""
    import torch

    from linalg_b200 import dev
    out = {}
    g = torch.Generator(device="cuda")
    g.manual_seed(20261108)
    A0 = torch.rand((n, n), dtype=torch.float64, device="cuda", generator=g)
    W = torch.empty_like(A0)
    ipiv = torch.zeros(n, dtype=torch.int32, device="cuda")
    info = torch.zeros(1, dtype=torch.int32, device="cuda")
    reps = max(2, min(steps, 3))
    ms = _best_ms(lambda: dev.getrf(n, n, W, n, ipiv, info), reps, setup=lambda: W.copy_(A0))
    out["dgetrf_tflops"] = (2.0 / 3.0) * n ** 3 / (ms * 1e-3) * 1e-12
    out["dgetrf_ms"] = ms
    out["dgetrf_info"] = int(info.item())
    # parity probe: ||P A x - L (U x)|| / (||A|| ||x||); torch views are the transposes of our
    # column-major matrices, so W.T is the matrix in math orientation
    xm = torch.rand(n, dtype=torch.float64, device="cuda", generator=g)
    LU = W.T
    Ux = torch.triu(LU) @ xm
    LUx = torch.tril(LU, -1) @ Ux + Ux
    Ax = A0.T @ xm
    piv = (ipiv.to(torch.int64) - 1).cpu().numpy()
    p = np.arange(n)
    for i in range(n):
        j = piv[i]
        if j != i:
            p[i], p[j] = p[j], p[i]
    PAx = Ax[torch.from_numpy(p).cuda()]
    out["dgetrf_probe_residual"] = float(((PAx - LUx).norm() / (A0.norm() * xm.norm())).item())

    # SPD matrix: symmetric, diagonally dominant (cheap variant)
    A0.add_(A0.T.clone()).mul_(0.5)
    A0.diagonal().add_(float(n))
    ms = _best_ms(lambda: dev.potrf(dev.LOWER, n, W, n, info), reps, setup=lambda: W.copy_(A0))
    out["dpotrf_tflops"] = (1.0 / 3.0) * n ** 3 / (ms * 1e-3) * 1e-12
    out["dpotrf_ms"] = ms
    out["dpotrf_info"] = int(info.item())
    Lm = torch.triu(W).T          # column-major lower triangle == upper triangle of the view
    Lx = Lm @ (Lm.T @ xm)
    Ax = A0.T @ xm
    out["dpotrf_probe_residual"] = float(((Ax - Lx).norm() / (A0.norm() * xm.norm())).item())
    del A0, W, LU, Lm
    torch.cuda.empty_cache()
    out["fp64_peak_tflops"] = FP64_DMMA_PEAK_TFLOPS
    out["dgetrf_frac_of_peak"] = out["dgetrf_tflops"] / FP64_DMMA_PEAK_TFLOPS
    out["dpotrf_frac_of_peak"] = out["dpotrf_tflops"] / FP64_DMMA_PEAK_TFLOPS
    out["dgetrf_roofline"] = _roofline_obj("dgetrf", out["dgetrf_tflops"], out["dgetrf_ms"],
                                           FP64_DMMA_PEAK_TFLOPS)
    out["dpotrf_roofline"] = _roofline_obj("dpotrf", out["dpotrf_tflops"], out["dpotrf_ms"],
                                           FP64_DMMA_PEAK_TFLOPS)

    # Float32 / Complex factorisations at the same N (flop counts: complex x4)
    for dt, let, fl in ((torch.float32, "s", 1.0), (torch.complex128, "z", 4.0)):
        A0 = torch.rand((n, n), dtype=dt, device="cuda", generator=g)
        W = torch.empty_like(A0)
        ms = _best_ms(lambda: dev.getrf(n, n, W, n, ipiv, info), 1, setup=lambda: W.copy_(A0))
        out[f"{let}getrf_ms"] = ms
        out[f"{let}getrf_tflops"] = fl * (2.0 / 3.0) * n ** 3 / (ms * 1e-3) * 1e-12
        out[f"{let}getrf_info"] = int(info.item())
        A0.add_(A0.T.conj().clone()).mul_(0.5)
        A0.diagonal().add_(float(n))
        ms = _best_ms(lambda: dev.potrf(dev.LOWER, n, W, n, info), 1, setup=lambda: W.copy_(A0))
        out[f"{let}potrf_ms"] = ms
        out[f"{let}potrf_tflops"] = fl * (1.0 / 3.0) * n ** 3 / (ms * 1e-3) * 1e-12
        out[f"{let}potrf_info"] = int(info.item())
        x = torch.rand(n, dtype=dt, device="cuda", generator=g)
        Lm = torch.triu(W).T
        res = (A0.T @ x - Lm @ (Lm.conj().T @ x)).norm() / (A0.norm() * x.norm())
        out[f"{let}potrf_probe_residual"] = float(res.item())
        del A0, W, Lm
        torch.cuda.empty_cache()
    return out


def time_other_types():
    """Float32 (3xTF32 tcgen05) and complex128 (DMMA) GEMM and the tall-skinny least squares of
    BASELINE config c5 (Float64 and Complex), device resident, one size each."""
    import torch

    from linalg_b200 import dev, ffi
    out = {}
    for dt, name, n, fl, want in ((torch.float32, "sgemm", 16384, 2, 3), (torch.complex128, "zgemm", 8192, 8, 2)):
        A = torch.rand((n, n), dtype=dt, device="cuda")
        B = torch.rand((n, n), dtype=dt, device="cuda")
        Cm = torch.zeros((n, n), dtype=dt, device="cuda")
        ms = _best_ms(lambda: dev.gemm(ffi.NO_TRANS, ffi.NO_TRANS, n, n, n, 1.0, A, n, B, n, 0.0, Cm, n))
        assert dev.last_gemm_path() == want, f"{name}: tensor-core path not taken"
        out[f"{name}_n"] = n
        out[f"{name}_tflops"] = fl * n ** 3 / (ms * 1e-3) * 1e-12
        # three full columns of C against a product formed in double precision
        cols = [0, n // 2 + 1, n - 1]
        hi = torch.complex128 if dt.is_complex else torch.float64
        ref = torch.stack([A.T.to(hi) @ B[c].to(hi) for c in cols]) if dt != torch.float32 else None
        if ref is None:
            Ad = A.T.to(torch.float64)
            ref = torch.stack([Ad @ B[c].to(torch.float64) for c in cols])
            del Ad
        got = Cm[cols].to(hi)
        out[f"{name}_relerr_vs_f64"] = float(((got - ref).norm() / ref.norm()).item())
        out[f"{name}_max_abs_err_vs_f64"] = float((got - ref).abs().max().item())
        del A, B, Cm, ref, got
        torch.cuda.empty_cache()
    out["tf32_peak_tflops_burst"] = TF32_PEAK_BURST_TFLOPS
    out["tf32_peak_tflops_sustained"] = TF32_PEAK_SUSTAINED_TFLOPS
    # the tensor pipe executes three TF32 products per useful one
    out["sgemm_frac_of_measured_tf32_peak"] = 3.0 * out["sgemm_tflops"] / TF32_PEAK_BURST_TFLOPS
    out["sgemm_frac_of_sustained_tf32_peak"] = 3.0 * out["sgemm_tflops"] / TF32_PEAK_SUSTAINED_TFLOPS
    m, n = 1 << 20, 512
    for dt, let, fl in ((torch.float64, "d", 1.0), (torch.complex128, "z", 4.0)):
        A = torch.rand((n, m), dtype=dt, device="cuda")
        b = torch.rand((1, m), dtype=dt, device="cuda")
        W, X = torch.empty_like(A), torch.empty_like(b)
        info = torch.zeros(1, dtype=torch.int32, device="cuda")
        ms = _best_ms(lambda: dev.gels(ffi.NO_TRANS, m, n, 1, W, m, X, m, info), 2,
                      setup=lambda: (W.copy_(A), X.copy_(b)))
        r = A.T @ X[0, :n] - b[0]
        out[f"{let}gels_1Mx512_ms"] = ms
        out[f"{let}gels_1Mx512_tflops"] = fl * (2.0 * m * n * n - (2.0 / 3.0) * n ** 3) / (ms * 1e-3) * 1e-12
        out[f"{let}gels_1Mx512_normal_eq_residual"] = float(
            ((A.conj() @ r).norm() / (A.norm() * r.norm())).item())
        del A, b, W, X
        torch.cuda.empty_cache()
    out.update(time_expm())
    return out


def time_expm(n=4096):
    """A consumer of the path that is a pure chain of products and one solve:
    the reference's expm (src/linalg/expm.cr) on device-resident operands, host matrix in, host
    matrix out; checked through exp(A) exp(-A) = I."""
    import time

    import numpy as np

    from linalg_b200 import la
    rng = np.random.default_rng(20261110)
    A = la.GeneralMatrix(np.asfortranarray((rng.random((n, n)) - 0.5) * (8.0 / np.sqrt(n))))
    A.expm()                                           # warm-up
    t0 = time.perf_counter()
    E = A.expm()
    ms = (time.perf_counter() - t0) * 1e3
    Em = la.GeneralMatrix(-A.raw).expm()
    err = float(np.abs((E * Em).raw - np.eye(n)).max())
    return {"expm_n": n, "expm_ms_host_in_host_out": ms, "expm_times_expm_of_minus_a_minus_identity": err}


def dist_legs(world, rank):
    """Factorisations at N = 65536 and the TSQR least squares across the ranks of this job."""
    import torch.distributed as dist

    from linalg_b200 import dist as ladist
    from linalg_b200 import dist_measure as dm
    out = {}
    single = {}
    if rank == 0:
        for kind in ("potrf", "getrf"):
            single[kind] = dm.run_single_gpu(kind, DIST_N)
    if world == 1:
        for kind in ("potrf", "getrf"):
            s = single[kind]
            out.update({f"dist_{kind}_n": DIST_N, f"dist_{kind}_ms": s["ms"], f"dist_{kind}_tflops": s["tflops"],
                        f"dist_{kind}_speedup_vs_1gpu": 1.0, f"dist_{kind}_residual": s["residual"],
                        f"dist_{kind}_info": s["info"], f"dist_{kind}_grid": "1x1 (single-GPU routine)"})
        return out
    dist.barrier()
    grid = ladist.Grid(comm_ctas=DIST_COMM_CTAS)
    d = dm.run_potrf(DIST_N, DIST_NB, grid, reserve_sms=DIST_RESERVE_SMS, reps=2,
                     warm_n=DIST_NB * grid.Q * 2)
    lu_grid = grid if grid.P == 1 else ladist.Grid(1, world, comm_ctas=DIST_COMM_CTAS)
    e = dm.run_getrf(DIST_N, DIST_NB, lu_grid, reserve_sms=DIST_RESERVE_SMS, reps=2,
                     warm_n=DIST_NB * world * 2)
    t = dm.run_tsqr(1 << 20, 512)
    for kind, r in (("potrf", d), ("getrf", e)):
        out.update({f"dist_{kind}_n": DIST_N, f"dist_{kind}_nb": DIST_NB, f"dist_{kind}_grid": r["grid"],
                    f"dist_{kind}_ms": r["ms"], f"dist_{kind}_tflops": r["tflops"],
                    f"dist_{kind}_residual": r["residual"], f"dist_{kind}_info": r["info"],
                    f"dist_{kind}_solve_ms": r["solve_ms"]})
        out["dist_potrf_solve_max_rel_err" if kind == "potrf" else "dist_getrf_solve_residual"] = \
            r["solve_max_rel_err" if kind == "potrf" else "solve_residual"]
        if rank == 0:
            out[f"dist_{kind}_1gpu_ms"] = single[kind]["ms"]
            out[f"dist_{kind}_1gpu_residual"] = single[kind]["residual"]
            out[f"dist_{kind}_speedup_vs_1gpu"] = single[kind]["ms"] / r["ms"]
    out.update({"dist_tsqr_m": t["m"], "dist_tsqr_n": t["n"], "dist_tsqr_ms": t["ms"],
                "dist_tsqr_tflops": t["tflops"], "dist_tsqr_max_rel_err": t["max_rel_err"],
                "dist_reserve_sms": DIST_RESERVE_SMS, "dist_comm_ctas": DIST_COMM_CTAS})
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist

    import linalg_b200
    from linalg_b200 import dev, ffi
    from linalg_b200 import dist as ladist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n = args.n
    c0, c1 = ladist.shard_columns(n, world, rank)   # this rank's block columns
    nc = c1 - c0
    g = torch.Generator(device="cuda")
    g.manual_seed(20261106)
    A = torch.rand((n, n), dtype=torch.float64, device="cuda", generator=g)    # col-major n x n
    g.manual_seed(20261107 + rank)
    B = torch.rand((nc, n), dtype=torch.float64, device="cuda", generator=g)   # n x nc block
    Cm = torch.zeros((nc, n), dtype=torch.float64, device="cuda")
    lib = linalg_b200.lib()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(x):
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for _ in range(args.warmup):
        Cm.zero_()
        dev.gemm(ffi.NO_TRANS, ffi.NO_TRANS, n, nc, n, 1.0, A, n, B, n, 1.0, Cm, n)
    assert dev.last_gemm_path() == 1, "DMMA tensor-core path was not taken"
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
           for _ in range(args.steps)]
    with ClockSampler(local) as clk:
        e0.record()
        for k0, k1 in kev:
            Cm.zero_()
            k0.record()
            dev.gemm(ffi.NO_TRANS, ffi.NO_TRANS, n, nc, n, 1.0, A, n, B, n, 1.0, Cm, n)
            k1.record()
        e1.record()
        barrier()
    ms = allmax(e0.elapsed_time(e1) / args.steps)
    value = 2.0 * n ** 3 / (ms * 1e-3) * 1e-12
    # one dev.gemm call = `launches` launches of the same kernel (k-chunks of 4096 when k >= 8192)
    launches = -(-n // DGEMM_KCHUNK) if (n >= 2 * DGEMM_KCHUNK and n * nc >= 4096 * 4096) else 1
    kern_ms = float(np.mean([a.elapsed_time(b) for a, b in kev])) / launches
    achieved = 2.0 * n * nc * (n / launches) / (kern_ms * 1e-3) * 1e-12

    # spot check: three columns of C against numpy in float64
    cols = [0, nc // 2, nc - 1]
    ref = A.T.cpu().numpy() @ B[cols].T.cpu().numpy()
    got = Cm[cols].T.cpu().numpy()
    relerr = float(np.linalg.norm(got - ref) / np.linalg.norm(ref))
    assert relerr < 1e-12, f"dgemm residual {relerr}"
    extra = {"dgemm_relerr": relerr}

    # ---- e2e with host buffers
    e2e_steps = max(1, min(args.steps, 3))
    hA = torch.empty((n, n), dtype=torch.float64).pin_memory()
    hB = torch.empty((nc, n), dtype=torch.float64).pin_memory()
    hA.copy_(A)
    hB.copy_(B)
    torch.cuda.synchronize()
    del Cm, A, B
    torch.cuda.empty_cache()
    # Matrix#* multiplies into a freshly zeroed result (mult.cr:118-119): one zeroed C per step,
    # prepared before the clock starts (inputs are ready when a step begins)
    hCs = [torch.zeros((nc, n), dtype=torch.float64).pin_memory() for _ in range(e2e_steps)]
    if world == 1:
        f = lib.cblas_dgemm
        f.argtypes = [ctypes.c_int] * 6 + [ctypes.c_double, ctypes.c_void_p, ctypes.c_int,
                                           ctypes.c_void_p, ctypes.c_int, ctypes.c_double,
                                           ctypes.c_void_p, ctypes.c_int]
        f.restype = None

        def host_gemm(pa, pb, pc):
            f(ffi.COL_MAJOR, ffi.NO_TRANS, ffi.NO_TRANS, n, nc, n, 1.0, pa, n, pb, n, 1.0, pc, n)

        def e2e_step(hc):
            host_gemm(hA.data_ptr(), hB.data_ptr(), hc.data_ptr())
        e2e_how = "cblas_dgemm (drop-in symbol), pinned host buffers"
        h2d = 8 * (n * n + 2 * n * nc)
    else:
        plan = ladist.HostGemmPlan(n, nc)
        a0, a1 = plan.a_slice()
        hA_slice = hA[a0:a1]

        def e2e_step(hc):
            plan.run(hA_slice, hB, hc)
        e2e_how = ("dist.HostGemmPlan: 1/N of A + block columns of B, C per rank from pinned host "
                   "memory, A all-gathered over NVLink")
        h2d = 8 * ((a1 - a0) * n + 2 * n * nc)
    warm = torch.zeros((nc, n), dtype=torch.float64).pin_memory()
    e2e_step(warm)   # warm-up (sizes the device workspace)
    del warm
    barrier()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        e2e_step(hCs[i])
    barrier()
    e2e_s = allmax((time.perf_counter() - t0) / e2e_steps)
    got = hCs[-1][cols].T.numpy()
    relerr2 = float(np.linalg.norm(got - ref) / np.linalg.norm(ref))
    assert relerr2 < 1e-12, f"host-buffer dgemm residual {relerr2}"
    e2e = {"value": 2.0 * n ** 3 / e2e_s * 1e-12, "unit": "TFLOP/s",
           "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 8 * n * nc,
           "ms_per_step": e2e_s * 1e3, "steps": e2e_steps, "how": e2e_how}
    del hCs
    if world == 1:
        # the same call on pageable memory (plain numpy arrays), as a Crystal caller's GC-heap
        # slices are: the shim stages through its pinned ring
        pA, pB = np.empty((n, n)), np.empty((nc, n))
        pA[:] = hA.numpy()
        pB[:] = hB.numpy()
        pCs = [np.zeros((nc, n)) for _ in range(e2e_steps + 1)]
        host_gemm(pA.ctypes.data, pB.ctypes.data, pCs[-1].ctypes.data)
        t0 = time.perf_counter()
        for i in range(e2e_steps):
            host_gemm(pA.ctypes.data, pB.ctypes.data, pCs[i].ctypes.data)
        pg_s = (time.perf_counter() - t0) / e2e_steps
        relerr3 = float(np.linalg.norm(pCs[0][cols].T - ref) / np.linalg.norm(ref))
        assert relerr3 < 1e-12, f"pageable cblas_dgemm residual {relerr3}"
        extra["e2e_pageable"] = {"value": 2.0 * n ** 3 / pg_s * 1e-12, "unit": "TFLOP/s",
                                 "ms_per_step": pg_s * 1e3, "steps": e2e_steps,
                                 "how": "cblas_dgemm on unpinned numpy buffers (the reference's "
                                        "GC-heap case)"}
        del pA, pB, pCs
    del hA, hB

    extra.update(dist_legs(world, rank))
    cpu = None
    if rank == 0:
        extra.update(time_factorizations(n, args.steps))
        extra.update(time_other_types())
        if world == 1:
            tf, cores, cfg, runs = cpu_dgemm_tflops(n, reps=1)
            cpu = {"value": tf, "unit": "TFLOP/s", "cores": cores, "kind": "reference",
                   "sample": f"OpenBLAS dgemm N={n} (the library the reference links, the full "
                             f"workload), best of {runs}, ~{2.0 * n ** 3 / tf * 1e-12:.1f} s each; {cfg}"}
        out = {
            "metric": "fp64_gemm_tflops", "value": value, "unit": "TFLOP/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": WORKLOAD.format(n=n),
                       "shard": f"block columns of B and C over {world} rank(s), A replicated",
                       "l2": "operands (3 x 2 GiB) exceed L2; no flush needed"},
            "e2e": e2e, "gpu_launches": args.steps * launches, "clocks": clk.summary(),
            "roofline": {"bound": "tensor", "achieved": achieved, "peak": FP64_DMMA_PEAK_TFLOPS,
                         "unit": "TFLOP/s", "frac": achieved / FP64_DMMA_PEAK_TFLOPS,
                         "traffic": DGEMM_DRAM_BYTES_PER_LAUNCH if (world == 1 and n == 16384) else None,
                         "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu "
                                           "metrics pass (profiles/r1e_dgemm16384_chunk_launches.csv); "
                                           "algorithmic bytes per launch 3.22e9 (A, B chunk once, C read+write)",
                         "kernel": "dgemm_dmma_kernel<A_MN=1,B_MN=0>",
                         "kernel_ms": kern_ms, "launches_per_step": launches,
                         "peak_source": "FP64 DMMA (mma.sync m8n8k4) register-resident peak measured "
                                        "by tools/microbench/peak_fp64.cu on this pool; "
                                        "MEASURED_PEAKS.json has no FP64 figure; nominal 37.2 at 1965 MHz",
                         "hbm_gbs_measured": _peaks().get("hbm_gbs")},
            "extra": extra,
        }
        if cpu is not None:
            out["cpu_baseline"] = cpu
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", type=int, default=16384)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
