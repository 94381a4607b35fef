#!/usr/bin/env python
"""How much device-to-host bandwidth does the box give when 1, 2, 4, 8 GPUs copy at once?

Each rank copies the C4 bitmap (10 000 rows x 6 272 B = 62.7 MB) from its GPU into its own pinned host buffer with one
plain cudaMemcpyAsync per copy, all ranks at the same time.  Prints per-rank and aggregate GB/s.  This is the
denominator behind every "bitmap to the host" figure of bench.py (SURVEY.md 8e expected eight parallel PCIe paths).

    python tools/d2h_scaling.py                       # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29555 tools/d2h_scaling.py
"""
import json
import os
import time

import torch


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist_

        dist = dist_
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    nbytes = 10_000 * 6272
    d = torch.zeros(nbytes, dtype=torch.uint8, device="cuda")
    h = torch.zeros(nbytes, dtype=torch.uint8).pin_memory()
    for _ in range(3):
        h.copy_(d, non_blocking=True)
    torch.cuda.synchronize()
    res = []
    for rep in range(5):
        if dist:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(20):
            h.copy_(d, non_blocking=True)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / 20
        res.append(dt)
    best = min(res)
    mine = torch.tensor([nbytes / best / 1e9], device="cuda", dtype=torch.float64)
    allr = [torch.zeros_like(mine) for _ in range(world)]
    if dist:
        dist.all_gather(allr, mine)
    else:
        allr = [mine]
    if rank == 0:
        per = [float(x.item()) for x in allr]
        print(json.dumps({"gpus_copying_at_once": world, "bytes_per_copy": nbytes, "per_rank_gbs": per, "aggregate_gbs": sum(per),
                          "ms_per_copy_rank0": best * 1e3}))
    if dist:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
