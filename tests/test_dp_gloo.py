"""The N>1 host logic on CPU: two gloo ranks shard a request list, "serve" their shard, and
rank 0 reassembles the outputs by request id; the job's step time is the max over ranks."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, num_reqs, q):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from pkgload import load_pkg
    dp = load_pkg("dp")
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = dp.request_shard(num_reqs, rank, world)
    local = {r: [r, r * 10, rank] for r in mine}          # a stand-in for the generated token ids
    merged = dp.gather_outputs(local, num_reqs)
    slow = dp.max_over_ranks(1.0 + rank)                    # rank 1 is the slow one
    if rank == 0:
        q.put((merged, slow, mine))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("num_reqs", [1, 7, 128])
def test_two_ranks_shard_and_merge(num_reqs):
    world, port = 2, 29650 + num_reqs % 50
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, num_reqs, q)) for r in range(world)]
    for p in procs:
        p.start()
    merged, slow, mine0 = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert slow == 2.0
    assert mine0 == list(range(0, num_reqs, 2))
    assert len(merged) == num_reqs
    for rid, toks in enumerate(merged):
        assert toks[:2] == [rid, rid * 10] and toks[2] == rid % 2


def test_shards_are_a_partition():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from pkgload import load_pkg
    dp = load_pkg("dp")
    for n in (0, 1, 5, 138, 512):
        for world in (1, 2, 4, 8):
            seen = sorted(r for k in range(world) for r in dp.request_shard(n, k, world))
            assert seen == list(range(n))
