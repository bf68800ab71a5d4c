"""CPU tests (gloo, world_size 2 and 3) of the N > 1 path's host logic: contiguous record-aligned shards,
count reduction, ordered concatenation.  The per-shard "search" is done by the oracle here (tests may use
it; on the GPU box the same logic drives libfxgpu in bench.py and the CLI)."""
import os
import random
import sys

import pytest
import torch.multiprocessing as mp

import helpers

ROOT = helpers.ROOT


def _worker(rank, world, port, data, pattern, flags, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from fxtract_b200 import lib, shard
    from oracle import fxoracle
    shards = shard.byte_shards(data, world, lambda b, n: lib.find_record_cut(b[:n]))
    off, ln = shards[rank]
    mine = data[off:off + ln]
    out = fxoracle.run(mine, flags, pattern=pattern) if ln else b""
    cnt = int(fxoracle.run(mine, flags + "c", pattern=pattern)) if ln else 0
    total = shard.reduce_counts(cnt, dist)
    whole = shard.gather_in_rank_order(out, dist)
    q.put((rank, shards, total, whole))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_search_equals_single(built, world):
    from oracle import fxoracle
    rng = random.Random(world)
    data = helpers.make_fastq(rng, 700)
    pattern = helpers.pick_substring(rng, data, 5, 8)
    ctx = mp.get_context("spawn")
    for flags in ("", "v"):
        q = ctx.Queue()
        port = 29600 + world * 10 + len(flags) + (os.getpid() % 200)
        procs = [ctx.Process(target=_worker, args=(r, world, port, data, pattern, flags, q)) for r in range(world)]
        for p in procs:
            p.start()
        res = sorted(q.get(timeout=120) for _ in range(world))
        for p in procs:
            p.join(timeout=60)
            assert p.exitcode == 0
        shards = res[0][1]
        assert all(r[1] == shards for r in res)
        assert shards[0][0] == 0 and sum(l for _, l in shards) == len(data)
        for (o, l), (o2, _) in zip(shards, shards[1:]):
            assert o + l == o2
        for o, l in shards:
            assert l == 0 or (data[o:o + 1] == b"@" and data[o + l - 1:o + l] == b"\n" and data[o:o + l].count(b"\n") % 4 == 0)
        want_out = fxoracle.run(data, flags, pattern=pattern)
        want_cnt = int(fxoracle.run(data, flags + "c", pattern=pattern))
        assert all(r[2] == want_cnt for r in res)
        assert res[0][3] == want_out


def test_record_ranges():
    from fxtract_b200 import shard
    for n, w in ((10, 3), (100_000_000, 8), (5, 8), (0, 2)):
        rr = shard.record_ranges(n, w)
        assert len(rr) == w and rr[0][0] == 0 and sum(c for _, c in rr) == n
        assert all(a + c == b for (a, c), (b, _) in zip(rr, rr[1:]))
        assert max(c for _, c in rr) - min(c for _, c in rr) <= 1
