"""world_size-2 gloo test of the segment sharding: CPU only, no compute calls."""
import os
import socket

import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from whisper_rs_b200 import shard


def test_partition_covers_everything_once():
    for n in (0, 1, 7, 64, 120, 10000):
        for world in (1, 2, 3, 4, 8):
            r = shard.all_ranges(n, world)
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
            sizes = [hi - lo for lo, hi in r]
            assert max(sizes) - min(sizes) <= 1


def test_beam_groups_stay_on_one_rank():
    for world in (2, 4, 8):
        for rank in range(world):
            lo, hi = shard.row_range(32, 5, rank, world)
            assert lo % 5 == 0 and hi % 5 == 0


def test_bad_arguments():
    with pytest.raises(ValueError):
        shard.segment_range(4, 2, 2)
    with pytest.raises(ValueError):
        shard.segment_range(-1, 0, 1)


def _worker(rank, world, port, n, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = shard.segment_range(n, rank, world)
    local = [{"segment": s, "tokens": [s, s + 1], "rank": rank} for s in range(lo, hi)]
    got = shard.gather_results(local, n)
    q.put((rank, [g["segment"] for g in got], [g["rank"] for g in got]))
    dist.barrier()
    dist.destroy_process_group()


def test_gather_world2_gloo():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    n, world = 7, 2
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, segs, owners in res:
        assert segs == list(range(n))
        assert owners == [0, 0, 0, 0, 1, 1, 1]
