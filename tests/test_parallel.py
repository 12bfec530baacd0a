"""world_size-2 gloo test (CPU) of the video sharding and the token gather."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from echr_b200.parallel import gather_tokens, shard_videos


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
