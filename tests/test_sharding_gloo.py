"""CPU: the N>1 host logic (unit sharding, max-over-ranks timing, aggregate throughput) under a real 2-process
torch.distributed group on the gloo backend."""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import phantom_fhe_boot_b200  # noqa: F401  (registers the package)
    from phantom_fhe_boot_b200 import sharding
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mine = list(sharding.shard_units(17, rank, world))
        ms = sharding.max_over_ranks(10.0 + 5.0 * rank, dist)      # rank 1 is the slow one
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
        # the row exchange of the limb-sharded key switch: uneven contiguous blocks, each broadcast by its owner
        import torch
        ranges = sharding.limb_ranges(5, world)          # (0, 3), (3, 5)
        rows = torch.zeros((5, 4), dtype=torch.int64)
        lo, hi = ranges[rank]
        rows[lo:hi] = torch.arange(lo * 4, hi * 4).view(hi - lo, 4) + 1000 * (rank + 1)
        sharding.exchange_rows(dist, rows, ranges)
        q.put((rank, mine, ms, gathered, rows.tolist()))
    finally:
        dist.destroy_process_group()


def test_two_rank_sharding_and_timing():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, m0, ms0, g0, rows0), (r1, m1, ms1, g1, rows1) = res
    want = [[1000 * (1 if i < 3 else 2) + 4 * i + j for j in range(4)] for i in range(5)]
    assert rows0 == rows1 == want                                                # every rank holds every row
    assert sorted(m0 + m1) == list(range(17)) and abs(len(m0) - len(m1)) <= 1     # every unit exactly once
    assert ms0 == ms1 == 15.0                                                   # MAX over ranks on both
    assert g0 == g1 == [m0, m1]


def test_shard_units_properties(fhe):
    from phantom_fhe_boot_b200 import sharding
    for n in (0, 1, 7, 8, 64, 1000):
        for world in (1, 2, 3, 4, 8):
            parts = [list(sharding.shard_units(n, r, world)) for r in range(world)]
            assert sum(parts, []) == list(range(n))
            assert max(map(len, parts)) - min(map(len, parts)) <= 1
    with pytest.raises(ValueError):
        sharding.shard_units(4, 2, 2)
    lr = sharding.limb_ranges(40, 8)
    assert lr == [(5 * r, 5 * r + 5) for r in range(8)]
    assert sharding.limb_ranges(10, 3) == [(0, 4), (4, 7), (7, 10)]
    assert sharding.aggregate_throughput(1, 50.0, 8) == 160.0
    assert sharding.max_over_ranks(3.5) == 3.5
