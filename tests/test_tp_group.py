"""Host logic of the tensor-parallel server group (atx/tp_group.py) on CPU: the follower ranks, fed only the
request stream rank 0 broadcasts, must make exactly the calls rank 0 makes — same admissions, same chunking,
same block tables, same chained (async-scheduled) steps — including requests rejected at admission and clients
that leave early. Real scheduler, allocator and LLMEngine.step(); the runner is a double (tests/fakes.py)."""
import asyncio

import pytest

from atx.engine import AsyncEngine
from atx.scheduler import SamplingParams
from atx.tp_group import apply_messages
from tests.fakes import make_logic_engine


class _InProcessGroup:
    """TPGroup double: the 'follower' is a second logic engine in this process, driven synchronously by exactly
    what TPGroup.broadcast would put on the pipe (atx/tp_group.py:_follower)."""

    def __init__(self, follower):
        self.follower = follower
        self.sent = []

    def broadcast(self, msgs):
        item = ("step", list(msgs))
        self.sent.append(item)
        apply_messages(self.follower, item[1])
        if self.follower.has_work():
            self.follower.step()

    def shutdown(self):
        pass


@pytest.mark.parametrize("async_sched", [True, False])
def test_followers_mirror_rank0_call_for_call(async_sched):
    lead, follow = make_logic_engine(async_scheduling=async_sched), make_logic_engine(async_scheduling=async_sched)
    group = _InProcessGroup(follow)
    ae = AsyncEngine(lead, tp_group=group)
    shared = list(range(100, 140))

    async def one(i, n_new, leave_after=None):
        got = []
        params = SamplingParams(temperature=0.0, max_tokens=n_new, ignore_eos=True, stream=True)
        async for out in ae.generate(shared + [i] * (5 + 3 * i), params, f"r{i}"):
            got = out.token_ids
            if leave_after is not None and len(got) >= leave_after:
                break                      # client goes away: rank 0 queues an abort, followers must see it too
        return got

    async def bad():
        with pytest.raises(ValueError):
            async for _ in ae.generate([5] * 300, SamplingParams(max_tokens=4), "too-long"):   # > max_model_len
                pass

    async def main():
        first = await asyncio.gather(one(0, 9), one(1, 4), one(2, 12, leave_after=3), bad(), one(3, 6), one(4, 7))
        second = await asyncio.gather(one(5, 5), one(6, 3))
        return first, second

    try:
        first, second = asyncio.run(main())
    finally:
        ae.shutdown()
    assert [len(x) for x in (first[0], first[1], first[4], first[5])] == [9, 4, 6, 7]
    assert [len(x) for x in second] == [5, 3]
    assert lead.trace, "rank 0 ran no steps"
    assert follow.trace == lead.trace                      # every step: same chunks, same block tables, same chaining
    assert any(t[0] for t in lead.trace) == async_sched    # chained steps are mirrored too
    assert any(m[0] == "abort" for item in group.sent for m in item[1])
    assert not follow.has_work() and not lead.has_work()
    assert follow.allocator.num_used == lead.allocator.num_used


def _seed_in_child(rid, conn):
    from atx.kv_cache import BlockAllocator
    from atx.scheduler import SamplingParams, Sequence

    conn.send(Sequence(rid, [1, 2, 3], SamplingParams(temperature=0.2), BlockAllocator(4, 16, False)).seed)


def test_default_seed_is_the_same_in_every_spawned_process():
    """Tensor-parallel followers are `spawn`ed processes that build their own Sequence from the broadcast
    request (atx/tp_group.py:_follower). At the reference's temperature 0.2 (llm/serve_llm.py:521) every rank
    must draw the same Gumbel noise, so the seed of a request without one cannot come from Python's salted
    str hash: two fresh interpreters (different PYTHONHASHSEED each) and this one must agree."""
    import multiprocessing as mp
    import os

    from atx.kv_cache import BlockAllocator
    from atx.scheduler import SamplingParams, Sequence, default_seed

    rid = "a3f9c1d2"
    mine = Sequence(rid, [1, 2, 3], SamplingParams(temperature=0.2), BlockAllocator(4, 16, False)).seed
    assert mine == default_seed(rid) and 0 <= mine < 2 ** 63
    assert default_seed("a3f9c1d3") != mine
    assert Sequence(rid, [1], SamplingParams(seed=0), BlockAllocator(4, 16, False)).seed == 0   # an explicit 0 is kept
    ctx = mp.get_context("spawn")
    seen = []
    for salt in ("1", "2"):
        os.environ["PYTHONHASHSEED"] = salt
        try:
            parent, child = ctx.Pipe()
            p = ctx.Process(target=_seed_in_child, args=(rid, child))
            p.start()
            seen.append(parent.recv())
            p.join(30)
        finally:
            os.environ.pop("PYTHONHASHSEED", None)
    assert seen == [mine, mine]
    assert default_seed(rid) == 0x7FFFFFFFFFFFFFFF & int.from_bytes(
        __import__("hashlib").blake2b(rid.encode(), digest_size=8).digest(), "little")


def _exit_at_once(conn):
    conn.close()


def test_rank0_notices_a_dead_or_failed_follower():
    """broadcast() runs before every engine-loop iteration of rank 0 (atx/engine.py:_run): a follower that has
    exited or reported a failure must surface as an error there, not as a step whose peers never answer."""
    import multiprocessing as mp

    from atx.tp_group import TPGroup

    ctx = mp.get_context("spawn")
    parent, child = ctx.Pipe(duplex=True)
    p = ctx.Process(target=_exit_at_once, args=(child,))
    p.start()
    p.join(30)
    child.close()
    g = object.__new__(TPGroup)
    g._conns, g._procs = [parent], [p]
    with pytest.raises(RuntimeError, match="rank 1"):
        g.broadcast([("abort", "x")])

    class _Alive:
        exitcode = None

        def is_alive(self):
            return True

    a, b = ctx.Pipe(duplex=True)
    g._conns, g._procs = [a], [_Alive()]
    g.broadcast([("abort", "x")])
    assert b.recv() == ("step", [("abort", "x")])
    b.send(("fatal", 1, "RuntimeError('device fault')"))
    with pytest.raises(RuntimeError, match="device fault"):
        g.broadcast([])
