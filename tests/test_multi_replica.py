"""N > 1 host logic on CPU: the data-parallel router / replica pool (real processes, fake engines) and
the cross-rank timing reduction of bench.py under torch.distributed `gloo`, world_size 2."""
import asyncio
import os
import socket

import pytest
import torch
import torch.multiprocessing as tmp

from atx.config import EngineConfig
from atx.router import ReplicaPool, Router
from atx.scheduler import SamplingParams
from tests.fakes import fake_replica_worker


def test_router_policy_prefix_affinity_then_least_loaded():
    r = Router(3, 16)
    shared = list(range(64))
    a = r.pick(shared + [1, 2, 3], 100)
    b = r.pick(shared + [9, 9], 100)
    assert a == b                                   # same first 4 blocks -> same replica (hot prefix cache)
    c = r.pick(list(range(500, 600)), 100)
    assert c != a                                   # unrelated prompt -> lightest replica
    d = r.pick([7] * 10, 100)                       # shorter than one block: no affinity key, least loaded
    assert d not in (a, c)
    for _ in range(20):                             # the home replica spills once it is clearly busier
        last = r.pick(shared + [4], 1000)
    assert last != a or r.load[a] <= 2.0 * min(r.load) + 8192
    assert max(r.load) <= 2.0 * min(r.load) + 8192 + 1000
    r.done(a, 100)
    assert all(x >= 0 for x in r.load)
    # N fan-out bursts arriving interleaved at N idle replicas: every burst stays together on its own replica
    r = Router(4, 16)
    bursts = [[100 * b + i for i in range(64)] for b in range(4)]
    where = {b: set() for b in range(4)}
    for i in range(6):
        for b in range(4):
            where[b].add(r.pick(bursts[b] + [i] * 200, 1280))
    assert all(len(v) == 1 for v in where.values()) and len({next(iter(v)) for v in where.values()}) == 4


def test_replica_pool_streams_routes_and_reports_errors():
    pool = ReplicaPool("fake", EngineConfig(block_size=16), 2, worker=fake_replica_worker)
    try:
        assert pool.kv_cache_info() == (201, 16) and pool.vocab_size == 1024

        async def one(rid, ids, n=3):
            outs = [o async for o in pool.generate(ids, SamplingParams(max_tokens=n), rid)]
            return outs

        async def scenario():
            shared = list(range(64))
            res = await asyncio.gather(one("a", shared + [1]), one("b", shared + [2]), one("c", list(range(200, 300))))
            return res
        res = asyncio.run(scenario())
        for outs in res:
            assert [len(o.token_ids) for o in outs] == [1, 2, 3] and outs[-1].finished
            assert outs[-1].finish_reason == "length" and not outs[0].finished
        rep = [outs[-1].token_ids[0] // 1000 for outs in res]
        assert rep[0] == rep[1] and rep[2] != rep[0]          # affinity + balance across real processes
        assert pool.router.load == [0, 0]                      # costs returned on completion
        with pytest.raises(ValueError, match="too long"):
            asyncio.run(one("big", list(range(600))))
        with pytest.raises(ValueError, match="duplicate"):
            async def dup():
                g1 = pool.generate([1, 2, 3], SamplingParams(max_tokens=2), "same")
                await g1.__anext__()
                async for _ in pool.generate([1, 2, 3], SamplingParams(max_tokens=2), "same"):
                    pass
            asyncio.run(dup())
    finally:
        pool.shutdown()


def test_replica_failure_fails_waiters_loudly():
    pool = ReplicaPool("explode", EngineConfig(block_size=16), 1, worker=fake_replica_worker)
    try:
        async def go():
            async for _ in pool.generate([1, 2, 3], SamplingParams(max_tokens=2), "x"):
                pass
        with pytest.raises(RuntimeError, match="replica 0 failed"):
            asyncio.run(go())
    finally:
        pool.shutdown()


@pytest.mark.parametrize("how", ["explode1", "vanish1"])
def test_dead_replica_is_taken_out_of_rotation(how):
    """Replica 1 dies on its first request (with a 'fatal' message, or silently: EOF on its pipe). Its request
    fails loudly, every later request is served by replica 0, and with no replica left generate() refuses at once
    (it used to keep routing to the dead one and wait forever)."""
    pool = ReplicaPool(how, EngineConfig(block_size=16), 2, worker=fake_replica_worker)
    seen = []
    pool.set_batch_size_observer(seen.append)
    try:
        async def one(rid, ids, n=2):
            return [o async for o in pool.generate(ids, SamplingParams(max_tokens=n), rid)]

        async def scenario():
            first = await asyncio.gather(one("a", [1] * 70), one("b", [2] * 70), return_exceptions=True)
            later = await asyncio.gather(*[one(f"c{i}", [3 + i] * 70) for i in range(4)])
            return first, later
        first, later = asyncio.run(scenario())
        failed = [r for r in first if isinstance(r, Exception)]
        served = [r for r in first if not isinstance(r, Exception)]
        assert len(failed) == 1 and "replica 1 failed" in str(failed[0]) and len(served) == 1
        assert all(outs[-1].finished and outs[-1].token_ids[0] // 1000 == 0 for outs in later)   # all on replica 0
        assert pool.router.alive == [True, False] and pool.router.load == [0, 0]
        assert seen and set(seen) == {1, 2}                     # llm_batch_size gets the replicas' step sizes
        assert pool.stats()["kv_used_blocks"] >= 7
        pool._replica_died(0, "test: the last one goes too")
        with pytest.raises(RuntimeError, match="no engine replica is alive"):
            asyncio.run(one("z", [9] * 70))
    finally:
        pool.shutdown()


def _rank_main(rank, world, port, q):
    import torch.distributed as dist

    import bench

    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    dist.barrier()
    mx = bench.max_over_ranks(dist, [10.0 + rank, 5.0 - rank], "cpu")
    allv = bench.gather_over_ranks(dist, [rank * 10 + 1, rank * 10 + 2], world)
    q.put((rank, mx, allv))
    dist.destroy_process_group()


def test_bench_cross_rank_reduction_gloo_world2():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = tmp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_rank_main, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, mx, allv in got:
        assert mx == [11.0, 5.0]                    # MAX over ranks: the slowest rank defines the time
        assert allv == [1, 2, 11, 12]               # every rank sees every rank's TTFT samples
    import bench

    assert bench.max_over_ranks(None, [3.0], "cpu") == [3.0] and bench.gather_over_ranks(None, [1], 1) == [1]
