#!/usr/bin/env python3
"""Host -> device bandwidth with N GPUs copying at once (run under torchrun, one rank per GPU): what bounds the end-to-end
numbers at N = 8 -- the per-GPU link, a shared PCIe switch, or host memory placement.  Every rank copies a page-locked
2 GB buffer to its GPU 6 times, all ranks between the same two barriers; once from a buffer allocated wherever the
process happened to run, once from a buffer allocated after the process moved to the GPU's NUMA node (first-touch
placement).  Rank 0 prints one JSON line."""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import bind_to_gpu_numa_node


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    nbytes = 2 << 30
    dev = torch.empty(nbytes, dtype=torch.uint8, device="cuda")

    def measure(buf):
        dev.copy_(buf, non_blocking=True)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(6):
            dev.copy_(buf, non_blocking=True)
        torch.cuda.synchronize()
        mine = 6 * nbytes / (time.perf_counter() - t0) / 1e9
        if world > 1:
            dist.barrier()
        t = torch.tensor([mine], dtype=torch.float64, device="cuda")
        if world > 1:
            allv = [torch.zeros_like(t) for _ in range(world)]
            dist.all_gather(allv, t)
            return [float(v.item()) for v in allv]
        return [mine]

    unbound = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    unbound.fill_(1)
    a = measure(unbound)
    del unbound
    numa = bind_to_gpu_numa_node(local)
    bound = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    bound.fill_(1)
    b = measure(bound)
    nodes = [None] * world
    if world > 1:
        dist.all_gather_object(nodes, numa)
    else:
        nodes = [numa]
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps({"probe": "pinned host -> device, %d GPUs at once, 2 GiB x 6 per GPU" % world, "n_gpus": world,
                          "per_gpu_GBps_unbound": [round(v, 1) for v in a], "sum_GBps_unbound": round(sum(a), 1),
                          "per_gpu_GBps_numa_local": [round(v, 1) for v in b], "sum_GBps_numa_local": round(sum(b), 1),
                          "numa": nodes}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
