"""N > 1 host logic on CPU: world_size-2 gloo processes shard streams with no data-path collective
and aggregate throughput as (sum of audio, max of time)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from pocket_tts_b200.sharding import LeastLoadedPlacer, aggregate_throughput, shard_streams


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_streams, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = shard_streams(n_streams, world, rank)
    # every stream is owned by exactly one rank: gather the ownership lists and check the partition
    lists = [None] * world
    dist.all_gather_object(lists, mine)
    flat = sorted(i for l in lists for i in l)
    assert flat == list(range(n_streams))
    # per-rank "work": rank r is slower; job time is the max, audio is the sum
    audio, ms = aggregate_throughput(len(mine) * 0.08 * 10, 5.0 + rank)
    if rank == 0:
        out.put((audio, ms, [len(l) for l in lists]))
    dist.destroy_process_group()


def test_two_rank_sharding_and_aggregation():
    world, n_streams = 2, 129
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_streams, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    audio, ms, sizes = q.get(timeout=10)
    assert sizes == [65, 64]
    assert audio == pytest.approx(n_streams * 0.8) and ms == pytest.approx(6.0)


def test_round_robin_is_a_partition_for_ragged_counts():
    for n in (0, 1, 7, 64, 2048):
        for world in (1, 2, 4, 8):
            parts = [shard_streams(n, world, r) for r in range(world)]
            assert sorted(i for p in parts for i in p) == list(range(n))
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1
    with pytest.raises(ValueError):
        shard_streams(4, 2, 2)


def test_least_loaded_placer():
    pl = LeastLoadedPlacer(world_size=3, capacity=2)
    ranks = [pl.admit() for _ in range(6)]
    assert ranks == [0, 1, 2, 0, 1, 2]
    with pytest.raises(RuntimeError):
        pl.admit()
    pl.release(1)
    assert pl.admit() == 1
    with pytest.raises(ValueError):
        LeastLoadedPlacer(2, 1).release(0)


def test_single_process_aggregate_is_identity():
    assert aggregate_throughput(3.0, 7.0) == (3.0, 7.0)
