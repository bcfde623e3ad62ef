"""CPU test of the N>1 path: two ranks over gloo agree on the shard layout without exchanging image
data, cover the batch exactly once, and the timing reduction bench.py uses (max over ranks) works."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from nvimagecodec_b200 import sharding


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, sizes, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = sharding.shard_for_rank(sizes, rank, world)
    # no data-path collective: only the bookkeeping bench.py does (sum of work, max of time)
    work = torch.tensor([float(sum(sizes[i] for i in mine)), float(len(mine))], dtype=torch.float64)
    t = torch.tensor([1.0 + rank], dtype=torch.float64)
    dist.all_reduce(work, op=dist.ReduceOp.SUM)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    if rank == 0:
        np.save(os.path.join(out_dir, "result.npy"), np.array([work[0].item(), work[1].item(), t[0].item()]))
        flat = sorted(i for part in gathered for i in part)
        assert flat == list(range(len(sizes)))
        assert gathered == sharding.partition_by_bytes(sizes, world)
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_shard_by_compressed_bytes(tmp_path):
    rng = np.random.default_rng(3)
    sizes = [int(x) for x in rng.integers(20_000, 700_000, size=513)]
    port = _free_port()
    mp.spawn(_worker, args=(2, port, sizes, str(tmp_path)), nprocs=2, join=True)
    total, count, tmax = np.load(tmp_path / "result.npy")
    assert int(total) == sum(sizes) and int(count) == len(sizes) and tmax == 2.0
    parts = sharding.partition_by_bytes(sizes, 2)
    assert sharding.imbalance(sizes, parts) < 1.001
