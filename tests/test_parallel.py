"""world_size-2 gloo test (CPU) of the video sharding and the token gather."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from echr_b200.parallel import JobPlan, gather_tokens, shard_videos


def test_shard_videos_covers_everything_in_order():
    counts = [3, 5, 4, 100, 2, 2, 7, 1]
    for world in (1, 2, 3, 4, 8):
        spans = [shard_videos(counts, r, world) for r in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == len(counts)
        for a, b in zip(spans, spans[1:]):
            assert a[1] == b[0]
    lo, hi = shard_videos([4] * 64, 1, 2)
    assert (lo, hi) == (32, 64)


def _worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    counts = [3, 5, 4, 1, 2]
    lo, hi = shard_videos(counts, rank, world)
    n = sum(counts[lo:hi])
    first = sum(counts[:lo])
    tok = (torch.arange(first, first + n, dtype=torch.int32).unsqueeze(1) * 100 +
           torch.arange(30, dtype=torch.int32).unsqueeze(0))
    full = gather_tokens(tok)
    want = (torch.arange(sum(counts), dtype=torch.int32).unsqueeze(1) * 100 +
            torch.arange(30, dtype=torch.int32).unsqueeze(0))
    ret[rank] = bool(torch.equal(full, want))
    dist.destroy_process_group()


def test_gather_tokens_gloo_world2():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(2, port, ret), nprocs=2, join=True)
    assert ret[0] and ret[1]


def test_job_plan_is_a_partition_with_equal_batch_counts():
    import numpy as np
    rs = np.random.RandomState(0)
    counts = [int(c) for c in rs.randint(3, 6, size=4096)]
    for world, bv in ((1, 512), (2, 512), (3, 512), (8, 512), (8, 100), (5, 7)):
        plan = JobPlan(counts, world, bv)
        flat = [ab for r in range(world) for ab in plan.batches[r]]
        assert flat[0][0] == 0 and flat[-1][1] == len(counts)
        assert all(a[1] == b[0] for a, b in zip(flat, flat[1:]))
        assert all(len(plan.batches[r]) == plan.n_batches for r in range(world))
        assert max(b - a for a, b in flat) <= bv
        assert sum(sum(e) for e in plan.events) == sum(counts)
    tiny = JobPlan([4, 4, 4], 2, 1)  # rank 1 has one video, rank 0 two: rank 1's second slot is empty
    assert tiny.n_batches == 2 and any(a == b for a, b in tiny.batches[1])


def _plan_worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    counts = [3, 5, 4, 1, 2, 6, 2]
    plan = JobPlan(counts, world, 2)
    first_row = [sum(counts[:v]) for v in range(len(counts) + 1)]
    gathered = []
    for b in range(plan.n_batches):  # every rank takes part in every slot's gather, also with an empty batch
        a, e = plan.batches[rank][b]
        tok = (torch.arange(first_row[a], first_row[e], dtype=torch.int32).unsqueeze(1) * 100 +
               torch.arange(30, dtype=torch.int32).unsqueeze(0))
        gathered.append(gather_tokens(tok, counts=plan.counts(b)))
    full = plan.assemble(gathered)
    want = (torch.arange(sum(counts), dtype=torch.int32).unsqueeze(1) * 100 + torch.arange(30, dtype=torch.int32).unsqueeze(0))
    ret[rank] = bool(torch.equal(full, want))
    dist.destroy_process_group()


def test_job_plan_gather_per_batch_gloo_world2():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_plan_worker, args=(2, port, ret), nprocs=2, join=True)
    assert ret[0] and ret[1]
