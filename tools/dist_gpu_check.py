"""Multi-GPU check of linalg_b200/dist.py, launched as
   python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
       --master-port 29511 tools/dist_gpu_check.py [--size 16384] [--nb 512] [--what potrf,getrf,tsqr]
       [--reserve 0,8] [--ctas 0,8] [--grids 2x4,1x8] [--single]
Runs every (grid, reserve, ctas) combination in ONE process group and prints one JSON line per
run from rank 0: device time (max over ranks), TFLOP/s of the whole job and the distributed probe
residual.  --single also times the single-GPU library routine on rank 0 (the denominator of the
speed-up)."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from linalg_b200 import dist as ladist          # noqa: E402
from linalg_b200 import dist_measure as dm      # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=16384)
    ap.add_argument("--nb", default="512")
    ap.add_argument("--what", default="potrf,getrf,tsqr")
    ap.add_argument("--reserve", default="0")
    ap.add_argument("--ctas", default="0")
    ap.add_argument("--grids", default="")
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--single", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    what = args.what.split(",")

    def say(kind, d):
        if rank == 0:
            print(json.dumps({"kind": kind, "world": world, **d}), flush=True)

    grids = [tuple(int(x) for x in g.split("x")) for g in args.grids.split(",") if g] or \
        [ladist.default_grid_shape(world)]
    nbs = [int(x) for x in args.nb.split(",")]
    if args.single and rank == 0:
        for kind in ("potrf", "getrf"):
            if kind in what:
                say(f"single_{kind}", dm.run_single_gpu(kind, args.size))
    if world > 1:
        dist.barrier()
    for ctas in [int(x) for x in args.ctas.split(",")]:
        if "potrf" in what:
            for P, Q in grids:
                grid = ladist.Grid(P, Q, comm_ctas=ctas)
                for nb in nbs:
                    for rs in [int(x) for x in args.reserve.split(",")]:
                        d = dm.run_potrf(args.size, nb, grid, reserve_sms=rs, reps=args.reps,
                                         warm_n=nb * Q * 2)
                        say("potrf", {"comm_ctas": ctas, **d})
        if "getrf" in what:
            grid = ladist.Grid(1, world, comm_ctas=ctas)
            for nb in nbs:
                for rs in [int(x) for x in args.reserve.split(",")]:
                    d = dm.run_getrf(args.size, nb, grid, reserve_sms=rs, reps=args.reps,
                                     warm_n=nb * world * 2)
                    say("getrf", {"comm_ctas": ctas, **d})
    if "tsqr" in what:
        say("tsqr", dm.run_tsqr(1 << 20, 512))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
