"""Stream sharding + counter gather across ranks (world_size 2, gloo, CPU)."""
import os
import socket

import pytest

from m17_cxx_demod_b200.shard import gather_counters, owner_of, shard_range


def test_shard_ranges_partition_the_channels():
    for n in (1, 7, 1024, 65536, 1000003):
        for world in (1, 2, 3, 4, 8):
            seen = 0
            prev_hi = 0
            for r in range(world):
                lo, hi = shard_range(n, world, r)
                assert lo == prev_hi and lo <= hi
                for c in (lo, hi - 1):
                    if lo < hi:
                        assert owner_of(c, n, world) == r
                seen += hi - lo
                prev_hi = hi
            assert seen == n
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = shard_range(1000, world, rank)
    local = {"frames": 10 * (rank + 1), "samples": (hi - lo) * 48000}
    total = gather_counters(local, world)
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, total))


def test_counter_gather_world_size_2_gloo():
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert out[0] == out[1] == {"frames": 30, "samples": 1000 * 48000}
