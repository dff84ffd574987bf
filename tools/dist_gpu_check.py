"""Multi-GPU check of linalg_b200/dist.py, launched as
   python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
       --master-port 29511 tools/dist_gpu_check.py [n] [nb]
Each rank owns the block columns k % world == rank of a synthetic SPD matrix (built locally, no
rank ever holds the whole matrix); prints the device time (max over ranks) and a probe residual
||S x - L (L^T x)|| / (||S|| ||x||) computed with distributed matrix-vector products."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from linalg_b200 import dist as ladist  # noqa: E402


def local_columns(n, cols):
    """S[:, cols] of S = n*I + 0.5*sin(...) (symmetric by construction), as a (ncols, n) tensor."""
    i = torch.arange(n, device="cuda", dtype=torch.float64).view(1, n)
    j = cols.to(torch.float64).view(-1, 1)
    lo, hi = torch.minimum(i, j), torch.maximum(i, j)
    s = 0.5 * torch.sin(lo * 12.9898 + hi * 78.233 + 0.1)
    s[torch.arange(len(cols), device="cuda"), cols] += float(n)
    return s


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
    nb = int(sys.argv[2]) if len(sys.argv) > 2 else 512
    reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0))))
    nblk = (n + nb - 1) // nb
    cols = torch.cat([torch.arange(k * nb, min(n, (k + 1) * nb)) for k in range(nblk)
                      if ladist.owner_of_block(k, world) == rank]).cuda()
    S_loc = local_columns(n, cols)
    ops = ladist.DevOps()
    times = []
    for it in range(reps):
        a_local = S_loc.clone()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        info = ladist.potrf_block_cyclic(a_local, n, nb, ops)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        times.append(float(t.item()))
    # probe residual with distributed products (S symmetric: S x = sum_ranks S[:, cols] x[cols])
    g = torch.Generator(device="cuda").manual_seed(3)
    x = torch.rand(n, dtype=torch.float64, device="cuda", generator=g) - 0.5
    Sx = S_loc.T @ x[cols]
    # keep only the lower triangle of the owned columns: element (r, c) with r >= c
    rows = torch.arange(n, device="cuda").view(1, n)
    Lloc = torch.where(rows >= cols.view(-1, 1), a_local, torch.zeros((), dtype=torch.float64, device="cuda"))
    z = torch.zeros(n, dtype=torch.float64, device="cuda")
    z[cols] = Lloc @ x                               # (L^T x)[cols]
    if world > 1:
        dist.all_reduce(z)
        dist.all_reduce(Sx)
    LLx = Lloc.T @ z[cols]
    if world > 1:
        dist.all_reduce(LLx)
    if rank == 0:
        res = float(((Sx - LLx).abs().max() / (n * x.abs().sum())).item())
        ms = min(times)
        print(f"potrf_block_cyclic n={n} nb={nb} world={world}: info={info} {ms:.2f} ms "
              f"{n**3 / 3 / ms / 1e9:.2f} TFLOP/s (all ranks) probe residual={res:.2e}", flush=True)
    # TSQR
    m_loc, nn = 1 << 18, 256
    g2 = torch.Generator(device="cuda").manual_seed(11 + rank)
    A = torch.rand((nn, m_loc), dtype=torch.float64, device="cuda", generator=g2)
    xt = torch.rand((1, nn), dtype=torch.float64, device="cuda",
                    generator=torch.Generator(device="cuda").manual_seed(5))
    b = xt @ A                                   # consistent system: x = xt exactly
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    xs = ladist.lstsq_tsqr(A.clone(), b.clone(), ops)
    e1.record()
    torch.cuda.synchronize()
    if rank == 0:
        print(f"lstsq_tsqr rows/rank={m_loc} n={nn} world={world}: {e0.elapsed_time(e1):.2f} ms "
              f"err={float((xs - xt).abs().max().item()):.2e}", flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
