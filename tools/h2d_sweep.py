#!/usr/bin/env python
"""What the platform gives: plain pinned cudaMemcpyAsync host -> device, 1.6 GB per GPU (configs[1]'s split), all
ranks at once.  The ceiling hgm_map_split's end-to-end rate is measured against (bench.py `e2e`).

  python tools/h2d_sweep.py                                        # one GPU
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
      tools/h2d_sweep.py                                           # N GPUs, one process per GPU, concurrently

Per rank: one copy of the whole buffer, and the same bytes as 64 MiB pieces on one stream and round-robin on
three streams (the library's staging shape).  Rank 0 prints one JSON line: GB/s per GPU (min / max) and aggregate.
One cudaMemcpyAsync per piece; no batched-copy API.
"""
import json
import os
import time

import torch
import torch.distributed as dist


def main():
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    nbytes = 1_600_000_000
    host = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    host.fill_(rank + 1)
    dst = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    streams = [torch.cuda.Stream(device=dev) for _ in range(3)]
    piece = 64 << 20

    def whole():
        dst.copy_(host, non_blocking=True)

    def pieces(n_streams):
        for i, off in enumerate(range(0, nbytes, piece)):
            with torch.cuda.stream(streams[i % n_streams]):
                dst[off:off + piece].copy_(host[off:off + piece], non_blocking=True)

    def timed(fn, reps=5):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / reps
        gbs = torch.tensor([nbytes / dt / 1e9], dtype=torch.float64, device=dev)
        if world > 1:
            all_ = [torch.zeros_like(gbs) for _ in range(world)]
            dist.all_gather(all_, gbs)
            vals = [v.item() for v in all_]
        else:
            vals = [gbs.item()]
        return {"per_gpu_min": min(vals), "per_gpu_max": max(vals), "aggregate": sum(vals), "per_gpu": vals}

    out = {"n_gpus": world, "bytes_per_gpu": nbytes, "unit": "GB/s",
           "one_copy": timed(whole), "pieces_64MiB_1_stream": timed(lambda: pieces(1)),
           "pieces_64MiB_3_streams": timed(lambda: pieces(3))}
    if rank == 0:
        try:
            out["host_cores"] = len(os.sched_getaffinity(0))
            out["numa_nodes"] = len([d for d in os.listdir("/sys/devices/system/node") if d.startswith("node")])
        except OSError:
            pass
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
