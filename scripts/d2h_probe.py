"""Aggregate device->host copy rate of N GPUs copying at once into pinned host memory (one process per GPU under
torchrun): the platform ceiling of the host-to-host `e2e` number, which moves 1.59 GB of pixels per GPU per step.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 scripts/d2h_probe.py
"""
import json
import os
import time

import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
nbytes = 1592524800  # pixels of 256 x 1920x1080 RGB
src = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
dst = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
h2d_src = torch.empty(128 << 20, dtype=torch.uint8, pin_memory=True)
h2d_dst = torch.empty(128 << 20, dtype=torch.uint8, device="cuda")
res = {}
for name, fn, moved in (("d2h", lambda: dst.copy_(src, non_blocking=True), nbytes),
                        ("h2d_128MB", lambda: h2d_dst.copy_(h2d_src, non_blocking=True), 128 << 20)):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    res[name] = {"per_gpu_gbs": moved * 10 / float(dt) / 1e9, "aggregate_gbs": world * moved * 10 / float(dt) / 1e9}
if rank == 0:
    print(json.dumps({"n_gpus": world, "cpus": os.cpu_count(), **res}))
if world > 1:
    dist.destroy_process_group()
