#!/usr/bin/env python
"""Host <-> device copy ceiling of the box: what `e2e` can reach when N GPUs upload and download at once.
Run under torch.distributed.run with one rank per GPU.  For k = 1, 2, 4, ... N concurrently active ranks it times
pinned-memory cudaMemcpyAsync of 1 GiB host->device, device->host, and both directions at once (two streams),
and prints the per-GPU and aggregate GB/s.  Plain copies only: this is the machine's ceiling, not library code.
"""
import json
import os

import torch
import torch.distributed as dist


def main():
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    nbytes = 1 << 30
    h_in = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    h_in.fill_(1)
    d_a = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    d_b = torch.ones(nbytes, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    out = {}
    k = 1
    actives = []
    while k < world:
        actives.append(k)
        k *= 2
    actives.append(world)
    for active in actives:
        for mode in ("h2d", "d2h", "both"):
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            ms = 0.0
            if rank < active:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                reps = 4
                e0.record()
                for _ in range(reps):
                    if mode in ("h2d", "both"):
                        with torch.cuda.stream(s1):
                            d_a.copy_(h_in, non_blocking=True)
                    if mode in ("d2h", "both"):
                        with torch.cuda.stream(s2):
                            h_out.copy_(d_b, non_blocking=True)
                torch.cuda.current_stream().wait_stream(s1)
                torch.cuda.current_stream().wait_stream(s2)
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / reps
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            per_dir = nbytes / (float(t.item()) * 1e-3) / 1e9
            out["%d_active_%s" % (active, mode)] = {"ms_per_GiB": round(float(t.item()), 2), "GBs_per_gpu_per_direction": round(per_dir, 1),
                                                   "aggregate_GBs_per_direction": round(per_dir * active, 1)}
    if rank == 0:
        print(json.dumps({"world": world, "bytes_per_copy": nbytes, "results": out}, indent=1))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
