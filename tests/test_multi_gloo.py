"""The N>1 path on CPU: world_size-2 gloo processes exercise the stripe partition and the bench
bookkeeping (max elapsed over ranks, summed units). There is no data-path collective to test:
stripes are independent."""
import os
import socket
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_stripes, q):
    sys.path.insert(0, os.path.join(ROOT, "rust-hadoop-generated-by-llm_b200"))
    from gpu_ec import partition
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = partition.stripes_for_rank(n_stripes, rank, world)
    # every rank reports its stripes; the union must be a partition of range(n_stripes)
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    ms, units = partition.aggregate(10.0 + 5.0 * rank, float(len(mine)), dist)
    q.put((rank, gathered, ms, units))
    dist.barrier()
    dist.destroy_process_group()


def test_world2_partition_and_aggregate():
    world, n = 2, 37
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, q)) for r in range(world)]
    [p.start() for p in procs]
    res = [q.get(timeout=120) for _ in range(world)]
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    for rank, gathered, ms, units in res:
        flat = sorted(i for part in gathered for i in part)
        assert flat == list(range(n))
        assert all(i % world == r for r, part in enumerate(gathered) for i in part)  # stripe i -> rank i mod G
        assert ms == 15.0  # max over ranks
        assert units == float(n)  # summed
