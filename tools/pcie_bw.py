#!/usr/bin/env python
"""Host <-> device copy bandwidth of this box with every rank copying at once: the ceiling of the e2e (host-buffer)
number of bench.py.  Pinned host buffers, one H2D and one D2H stream per rank running concurrently, no kernels.

usage: python tools/pcie_bw.py                                     (one GPU)
       python -m torch.distributed.run --nproc-per-node 8 ... tools/pcie_bw.py   (all ranks at once)
Prints one JSON line from rank 0: per-rank and total GB/s in each direction."""
import json
import os
import time

import torch
import torch.distributed as dist


def main():
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl")
    nbytes = 1 << 30
    h_in = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True); h_in.fill_(1)
    h_out = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    d_in = torch.empty(nbytes, dtype=torch.uint8, device="cuda"); d_out = torch.ones(nbytes, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    res = {}
    for mode in ("h2d", "d2h", "both"):
        for rep in range(3):
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            t0 = time.perf_counter()
            n = 4
            for _ in range(n):
                if mode in ("h2d", "both"):
                    with torch.cuda.stream(s1):
                        d_in.copy_(h_in, non_blocking=True)
                if mode in ("d2h", "both"):
                    with torch.cuda.stream(s2):
                        h_out.copy_(d_out, non_blocking=True)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
        t = torch.tensor([dt], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        res[mode] = {"per_rank_gbs_each_direction": n * nbytes / float(t.item()) / 1e9,
                     "total_gbs_each_direction": world * n * nbytes / float(t.item()) / 1e9}
    if rank == 0:
        aff = sorted(os.sched_getaffinity(0))
        print(json.dumps({"ranks": world, "cpus": len(aff), **res}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
