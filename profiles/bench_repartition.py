#!/usr/bin/env python
"""Cost of one device-side re-partitioning at bench size (per rank: --bodies rods, scrambled ownership of a fraction).
usage: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 profiles/bench_repartition.py [--bodies 1000000]
Prints one JSON line on rank 0: wall ms of mb200_dist_repartition (max over ranks) when (a) the blocks are already curve
blocks (nothing much moves) and (b) 10 % of every rank's bodies belong elsewhere."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mundy_b200 import dist as mbd  # noqa: E402
from mundy_b200 import workloads as wl  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bodies", type=int, default=1_000_000)
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    dev = torch.device("cuda", torch.cuda.current_device())
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    comm = mbd.Comm()
    n = args.bodies * world
    w = wl.rods(n, phi=0.1)
    w = mbd.reorder(w, mbd.hilbert_order(w["center"]))
    lo, hi = mbd.owned_range(n, world, rank)

    def tens(ix):
        t = lambda x, d=torch.float64: torch.as_tensor(np.ascontiguousarray(x[ix]), dtype=d, device=dev)  # noqa: E731
        return t(w["center"]), t(w["radius"]), t(w["quat"]), t(w["length"]), t(w["kind"], torch.int32)

    def timed(ix):
        c, r, q, L, k = tens(ix)
        extra = torch.zeros((len(ix), 6), dtype=torch.float64, device=dev)
        best = None
        for _ in range(3):
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            t0 = time.perf_counter()
            res = mbd.repartition(comm, c, r, q, L, k, extra=extra, capacity=int(1.5 * args.bodies))
            torch.cuda.synchronize()
            ms = torch.tensor([1e3 * (time.perf_counter() - t0)], device=dev)
            if world > 1:
                dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            best = ms.item() if best is None else min(best, ms.item())
        return best, int(res["radius"].numel())

    own = np.arange(lo, hi)
    a_ms, a_n = timed(own)
    rng = np.random.default_rng(rank)
    # 10 % of the block swapped for bodies of the other ranks' blocks (a permutation of the global set keeps totals)
    perm = np.arange(n)
    rs = np.random.default_rng(7)
    pick = rs.choice(n, size=n // 10, replace=False)
    perm[pick] = pick[rs.permutation(len(pick))]
    b_ms, b_n = timed(perm[lo:hi])
    if rank == 0:
        print(json.dumps({"what": "mb200_dist_repartition wall ms (best of 3, max over ranks)", "n_gpus": world,
                          "bodies_per_gpu": args.bodies, "payload_doubles_per_body": 12 + 6,
                          "curve_blocks_ms": round(a_ms, 3), "curve_blocks_n_out_rank0": a_n,
                          "ten_percent_misplaced_ms": round(b_ms, 3), "ten_percent_misplaced_n_out_rank0": b_n}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
