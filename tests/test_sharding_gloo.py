"""N>1 host logic on CPU: two processes over gloo check that the shards of the ME work are disjoint, cover
everything, carry the references each rank needs, and that the throughput aggregation is max-time / sum-units."""
import importlib
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sh = importlib.import_module("svt-av1-1_b200.sharding")
    import bench
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    jobs = bench.gop_jobs()
    mine, need = sh.picture_shard(jobs, world, rank)
    rows = sh.sb_row_shard(34, world, rank)                       # 2160 / 64 -> 34 SB rows
    streams = sh.stream_shard(8, world, rank)
    gathered = [None] * world
    dist.all_gather_object(gathered, {"jobs": mine, "need": need, "rows": rows, "streams": streams})
    units, ms, ups = sh.aggregate(len(mine), 10.0 * (rank + 1), dist)
    if rank == 0:
        q.put((gathered, units, ms, ups))
    dist.barrier()
    dist.destroy_process_group()


def test_world_size_2_sharding():
    world, port = 2, 29533
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    gathered, units, ms, ups = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    sys.path.insert(0, ROOT)
    import bench
    all_jobs = bench.gop_jobs()
    got = [j for g in gathered for j in g["jobs"]]
    assert sorted(got) == sorted(all_jobs) and len(set(got)) == len(all_jobs)          # disjoint cover
    for g in gathered:                                                                   # references replicated
        for poc, _tl, r0, r1 in g["jobs"]:
            assert {poc, r0, r1} <= set(g["need"])
        assert abs(len(g["jobs"]) - len(all_jobs) / world) <= 8                          # balanced to one mini-GOP
    rows = sorted(g["rows"] for g in gathered)
    assert rows[0][0] == 0 and rows[0][0] + rows[0][1] == rows[1][0] and rows[1][0] + rows[1][1] == 34
    assert sorted(s for g in gathered for s in g["streams"]) == list(range(8))
    assert units == len(all_jobs) and ms == 20.0 and abs(ups - len(all_jobs) / 0.020) < 1e-6


@pytest.mark.parametrize("n,world", [(8, 3), (5, 8), (34, 4), (1, 2)])
def test_shard_helpers_cover(n, world):
    sys.path.insert(0, ROOT)
    sh = importlib.import_module("svt-av1-1_b200.sharding")
    assert sorted(s for r in range(world) for s in sh.stream_shard(n, world, r)) == list(range(n))
    spans = [sh.sb_row_shard(n, world, r) for r in range(world)]
    pos = 0
    for first, cnt in spans:
        assert first == pos
        pos += cnt
    assert pos == n
