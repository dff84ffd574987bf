"""Two-or-more-GPU check of linalg_b200/dist.py, launched as
   python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
       --master-port 29511 tools/dist_gpu_check.py [n] [nb]
Each rank owns the block columns k % world == rank; prints residual and device time."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from linalg_b200 import dist as ladist  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
    nb = int(sys.argv[2]) if len(sys.argv) > 2 else 512
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    if world > 1:
        dist.init_process_group("nccl")
    g = torch.Generator(device="cuda").manual_seed(11)
    G = torch.rand((n, n), dtype=torch.float64, device="cuda", generator=g) - 0.5
    S = G @ G.T / n + torch.eye(n, dtype=torch.float64, device="cuda")
    nblk = (n + nb - 1) // nb
    cols = torch.cat([torch.arange(k * nb, min(n, (k + 1) * nb)) for k in range(nblk)
                      if ladist.owner_of_block(k, world) == rank]).cuda()
    ops = ladist.DevOps()
    times = []
    for it in range(3):
        a_local = S[cols].contiguous()            # S symmetric: rows of S = columns of S
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
    # residual: gather L and compare L L^T with S on rank 0
    L = torch.zeros((n, n), dtype=torch.float64, device="cuda")
    L[cols] = a_local
    if world > 1:
        dist.all_reduce(L)
    if rank == 0:
        Lm = torch.tril(L.T)
        res = float(((Lm @ Lm.T - S).abs().max() / S.abs().max()).item())
        ms = min(times)
        print(f"potrf_block_cyclic n={n} nb={nb} world={world}: info={info} {ms:.2f} ms "
              f"{n**3 / 3 / ms / 1e9:.2f} TFLOP/s resid={res:.2e}", flush=True)
    # TSQR
    m_loc, nn = 1 << 18, 256
    A = torch.rand((nn, m_loc), dtype=torch.float64, device="cuda", generator=g)
    xt = torch.rand((1, nn), dtype=torch.float64, device="cuda",
                    generator=torch.Generator(device="cuda").manual_seed(5))
    b = xt @ A                                   # consistent system: x = xt exactly
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    x = ladist.lstsq_tsqr(A.clone(), b.clone(), ops)
    e1.record()
    torch.cuda.synchronize()
    if rank == 0:
        print(f"lstsq_tsqr rows/rank={m_loc} n={nn} world={world}: {e0.elapsed_time(e1):.2f} ms "
              f"err={float((x - xt).abs().max().item()):.2e}", flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
