#!/usr/bin/env python
"""bench.py — headline benchmark of the la_b200 backend (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--n 16384]

Metric: FP64 GEMM TFLOP/s at N = 16384 (the DGEMM leg of "DGEMM/DGETRF/DPOTRF TFLOP/s at
N=16384"), in the reference's `Matrix#*` call pattern: C <- 1.0*A*B + 1.0*C on a zeroed C,
column-major, NoTrans/NoTrans (src/linalg/mult.cr:26,64-69,119-120).  One step = one such
product.  DGETRF / DPOTRF at the same N are timed in the same run and reported under "extra".

  value   device-resident: operands already in HBM, CUDA events on the launching stream
          (la_b200_dgemm_dev), max over ranks.
  e2e     the drop-in symbol cblas_dgemm called with HOST pointers (pinned), i.e. H2D of A, B, C
          + kernel + D2H of C inside the timed region.
  N > 1   strong scaling: B and C are split by block columns over the ranks,
          A is replicated; no collective on the data path; value = 2N^3 / max-rank time.
  --impl reference   the reference's own arithmetic path — OpenBLAS dgemm through the same C ABI
          on the host cores (all threads) — on a bounded sample (N = 8192) of the workload.

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
CPU_SAMPLE_N = 8192
DGEMM_KCHUNK = 4096                     # the library runs long-k products as k-chunks (gemm_dmma.cu)
DGEMM_DRAM_BYTES_PER_LAUNCH = 18.82e9   # one k-chunk launch at N = 16384: 16.69 GB read + 2.14 GB written (ncu)
WORKLOAD = "dgemm N={n} NoTrans/NoTrans alpha=1 beta=1 on zeroed C (Matrix#*)"


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
    return 2.0 * n ** 3 / best * 1e-12, cores, cfg


def run_reference(args):
    """Reference arm: the reference's CPU arithmetic path on the box's host cores (rank 0)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    n = CPU_SAMPLE_N
    step, cores, cfg = _cpu_gemm_setup(n)
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = (time.perf_counter() - t0) / args.steps
    tf = 2.0 * n ** 3 / dt * 1e-12
    sample = f"dgemm N={n} (2N^3 = {2.0 * n ** 3:.3e} flop per step) of the N={args.n} workload"
    print(json.dumps({
        "impl": "reference", "metric": "fp64_gemm_tflops", "value": tf, "unit": "TFLOP/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": WORKLOAD.format(n=args.n), "cpu_sample": sample},
        "cpu_baseline": {"value": tf, "unit": "TFLOP/s", "cores": cores, "kind": "reference",
                         "sample": sample + "; " + cfg},
        "e2e": {"value": tf, "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def time_factorizations(n, steps):
    """DGETRF and DPOTRF at the same N, device resident, with probe-vector parity checks."""
    import torch

    from linalg_b200 import dev
    out = {}
    g = torch.Generator(device="cuda")
    g.manual_seed(20261108)
    A0 = torch.rand((n, n), dtype=torch.float64, device="cuda", generator=g)
    W = torch.empty_like(A0)
    ipiv = torch.zeros(n, dtype=torch.int32, device="cuda")
    info = torch.zeros(1, dtype=torch.int32, device="cuda")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def timed(fn, reps):
        best = 1e30
        for _ in range(reps):
            W.copy_(A0)
            torch.cuda.synchronize()
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        return best

    reps = max(2, min(steps, 3)) + 1
    ms = timed(lambda: dev.getrf(n, n, W, n, ipiv, info), reps)
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
    ms = timed(lambda: dev.potrf(dev.LOWER, n, W, n, info), reps)
    out["dpotrf_tflops"] = (1.0 / 3.0) * n ** 3 / (ms * 1e-3) * 1e-12
    out["dpotrf_ms"] = ms
    out["dpotrf_info"] = int(info.item())
    Lm = torch.triu(W).T          # column-major lower triangle == upper triangle of the view
    Lx = Lm @ (Lm.T @ xm)
    Ax = A0.T @ xm
    out["dpotrf_probe_residual"] = float(((Ax - Lx).norm() / (A0.norm() * xm.norm())).item())
    del A0, W
    torch.cuda.empty_cache()
    out.update(time_other_types())
    out["fp64_peak_tflops"] = FP64_DMMA_PEAK_TFLOPS
    out["dgetrf_frac_of_peak"] = out["dgetrf_tflops"] / FP64_DMMA_PEAK_TFLOPS
    out["dpotrf_frac_of_peak"] = out["dpotrf_tflops"] / FP64_DMMA_PEAK_TFLOPS
    return out


def time_other_types():
    """Float32 (3xTF32 tcgen05) and complex128 (DMMA) GEMM and the tall-skinny least squares of
    BASELINE config c5, device resident, one size each."""
    import torch

    from linalg_b200 import dev, ffi
    out = {}
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def best(fn, reps=3):
        b = 1e30
        for i in range(reps + 1):
            torch.cuda.synchronize()
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            if i:
                b = min(b, e0.elapsed_time(e1))
        return b

    for dt, name, n, fl, want in ((torch.float32, "sgemm", 16384, 2, 3), (torch.complex128, "zgemm", 8192, 8, 2)):
        A = torch.rand((n, n), dtype=dt, device="cuda")
        B = torch.rand((n, n), dtype=dt, device="cuda")
        Cm = torch.zeros((n, n), dtype=dt, device="cuda")
        ms = best(lambda: dev.gemm(ffi.NO_TRANS, ffi.NO_TRANS, n, n, n, 1.0, A, n, B, n, 0.0, Cm, n))
        assert dev.last_gemm_path() == want, f"{name}: tensor-core path not taken"
        # probe-vector check: C x == A (B x)
        x = torch.rand(n, dtype=dt, device="cuda")
        lhs = Cm.T @ x
        rhs = A.T @ (B.T @ x)
        out[f"{name}_n"] = n
        out[f"{name}_tflops"] = fl * n ** 3 / (ms * 1e-3) * 1e-12
        out[f"{name}_probe_relerr"] = float(((lhs - rhs).norm() / rhs.norm()).item())
        del A, B, Cm
        torch.cuda.empty_cache()
    m, n = 1 << 20, 512
    A = torch.rand((n, m), dtype=torch.float64, device="cuda")
    b = torch.rand((1, m), dtype=torch.float64, device="cuda")
    W, X = torch.empty_like(A), torch.empty_like(b)
    info = torch.zeros(1, dtype=torch.int32, device="cuda")

    def ls():
        W.copy_(A)
        X.copy_(b)
        dev.gels(ffi.NO_TRANS, m, n, 1, W, m, X, m, info)

    tcopy = best(lambda: (W.copy_(A), X.copy_(b)), reps=2)
    ms = best(ls, reps=2) - tcopy
    r = A.T @ X[0, :n] - b[0]
    out["dgels_1Mx512_ms"] = ms
    out["dgels_1Mx512_tflops"] = (2.0 * m * n * n - (2.0 / 3.0) * n ** 3) / (ms * 1e-3) * 1e-12
    out["dgels_1Mx512_normal_eq_residual"] = float(((A @ r).norm() / (A.norm() * r.norm())).item())
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist

    import linalg_b200
    from linalg_b200 import dev, ffi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n = args.n
    c0, c1 = (n * rank) // world, (n * (rank + 1)) // world   # this rank's block columns
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

    # ---- e2e through the drop-in symbol with pinned host buffers
    e2e_steps = max(1, min(args.steps, 2))
    hA = torch.empty((n, n), dtype=torch.float64).pin_memory()
    hB = torch.empty((nc, n), dtype=torch.float64).pin_memory()
    hC = torch.zeros((nc, n), dtype=torch.float64).pin_memory()
    hA.copy_(A)
    hB.copy_(B)
    torch.cuda.synchronize()
    del Cm, A, B
    torch.cuda.empty_cache()
    f = lib.cblas_dgemm
    f.argtypes = [ctypes.c_int] * 6 + [ctypes.c_double, ctypes.c_void_p, ctypes.c_int,
                                       ctypes.c_void_p, ctypes.c_int, ctypes.c_double,
                                       ctypes.c_void_p, ctypes.c_int]
    f.restype = None

    def e2e_step(hc):
        f(ffi.COL_MAJOR, ffi.NO_TRANS, ffi.NO_TRANS, n, nc, n, 1.0, hA.data_ptr(), n, hB.data_ptr(),
          n, 1.0, hc.data_ptr(), n)

    # Matrix#* multiplies into a freshly zeroed result (mult.cr:118-119): one zeroed pinned C per
    # step, prepared before the clock starts (inputs are ready when a step begins)
    hCs = [hC] + [torch.zeros((nc, n), dtype=torch.float64).pin_memory() for _ in range(e2e_steps - 1)]
    e2e_step(hC)   # warm-up (sizes the shim's device workspace)
    hC.zero_()
    barrier()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        e2e_step(hCs[i])
    barrier()
    e2e_s = allmax((time.perf_counter() - t0) / e2e_steps)
    hC = hCs[-1]
    got = hC[cols].T.numpy()
    relerr2 = float(np.linalg.norm(got - ref) / np.linalg.norm(ref))
    assert relerr2 < 1e-12, f"cblas_dgemm residual {relerr2}"
    e2e = {"value": 2.0 * n ** 3 / e2e_s * 1e-12, "unit": "TFLOP/s",
           "h2d_bytes_per_step": 8 * (n * n + 2 * n * nc), "d2h_bytes_per_step": 8 * n * nc,
           "ms_per_step": e2e_s * 1e3, "steps": e2e_steps}
    del hA, hB, hC, hCs

    cpu = None
    if rank == 0:
        extra.update(time_factorizations(n, args.steps))
        if world == 1:
            tf, cores, cfg = cpu_dgemm_tflops(CPU_SAMPLE_N)
            cpu = {"value": tf, "unit": "TFLOP/s", "cores": cores, "kind": "reference",
                   "sample": f"OpenBLAS dgemm N={CPU_SAMPLE_N} (the library the reference links), "
                             f"best of 3, ~{2.0 * CPU_SAMPLE_N ** 3 / tf * 1e-12:.1f} s each; {cfg}"}
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
