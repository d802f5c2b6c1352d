"""CPU, world_size 2 over gloo: the host-side logic of the multi-GPU path -- stream -> GPU
assignment and the max-over-ranks / aggregate-throughput reduction bench.py uses.  The data
path itself has no collective (independent streams), so there is nothing else to emulate."""
import os
import socket

import pytest

from nnstreamer_b200 import sharding


def test_stream_assignment_is_a_partition():
    for world in (1, 2, 4, 8):
        seen = []
        for r in range(world):
            mine = sharding.streams_of_rank(64, world, r)
            assert all(sharding.stream_to_gpu(s, world) == r for s in mine)
            assert len(mine) == 64 // world
            seen += mine
        assert sorted(seen) == list(range(64))
    with pytest.raises(ValueError):
        sharding.stream_to_gpu(0, 0)


def test_reductions_without_a_process_group():
    assert sharding.max_over_ranks(1.5, None) == 1.5
    assert sharding.aggregate_throughput(32, 0.5, None) == 64.0
    assert sharding.gather_over_ranks([3, 4.5], None) == [[3.0, 4.5]]


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mine = sharding.streams_of_rank(8, world, rank)
        # each rank pretends its step took (rank + 1) * 10 ms for len(mine) * 32 frames
        seconds = 0.010 * (rank + 1)
        slowest = sharding.max_over_ranks(seconds, dist)
        agg = sharding.aggregate_throughput(len(mine) * 32, seconds, dist)
        gathered = sharding.gather_over_ranks([rank, seconds], dist)
        assert gathered == [[0.0, 0.010], [1.0, 0.020]]
        sharding.barrier(dist)
        q.put((rank, mine, slowest, agg))
    finally:
        dist.destroy_process_group()


def test_two_ranks_over_gloo():
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[0][1] == [0, 2, 4, 6] and res[1][1] == [1, 3, 5, 7]
    for rank, mine, slowest, agg in res:
        assert abs(slowest - 0.020) < 1e-12                   # MAX over ranks
        assert abs(agg - (8 * 32) / 0.020) < 1e-6             # all units / slowest rank
