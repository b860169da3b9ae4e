#!/usr/bin/env python
"""Host-link probe for the multi-rank `e2e` figure: every rank copies 1 GiB of pinned host memory to its GPU and back at the same
time (plain cudaMemcpyAsync through torch), alone and all together.  Prints one JSON line from rank 0: per-rank and aggregate GB/s.
Run under torchrun like bench.py.  No library code involved: it measures the box, not the repo."""
import json
import os
import time

import torch
import torch.distributed as dist


def main():
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n = 1 << 27  # doubles = 1 GiB
    host = torch.empty(n, dtype=torch.float64).pin_memory()
    host.fill_(1.0)
    dev = torch.empty(n, dtype=torch.float64, device="cuda")

    def timed(fn, together):
        fn()
        torch.cuda.synchronize()
        if world > 1 and together:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        return 3 * n * 8 / (time.perf_counter() - t0) / 1e9

    out = {}
    for name, fn in (("h2d", lambda: dev.copy_(host, non_blocking=True)), ("d2h", lambda: host.copy_(dev, non_blocking=True))):
        # all ranks at once
        g = torch.tensor([timed(fn, True)], device="cuda", dtype=torch.float64)
        if world > 1:
            lo = g.clone()
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            dist.all_reduce(g, op=dist.ReduceOp.SUM)
            out[name + "_together_min_rank_GBs"] = round(float(lo.item()), 1)
        out[name + "_together_sum_GBs"] = round(float(g.item()), 1)
        # one rank at a time
        alone = 0.0
        for r in range(world):
            if world > 1:
                dist.barrier()
            if r == rank:
                alone = timed(fn, False)
            if world > 1:
                dist.barrier()
        a = torch.tensor([alone], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(a, op=dist.ReduceOp.MIN)
        out[name + "_alone_min_rank_GBs"] = round(float(a.item()), 1)
    if rank == 0:
        print(json.dumps({"probe": "pinned host <-> device, 1 GiB x 3 per rank", "ranks": world, **out}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
