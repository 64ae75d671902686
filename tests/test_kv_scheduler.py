"""Host logic on CPU: block allocator (bit-exact against oracle/kv_ref.py), prefix cache, and the
continuous-batching scheduler driven by a fake forward pass."""
import random

import pytest

from atx.kv_cache import BlockAllocator, OutOfBlocks, SequenceBlocks, block_hash, kv_capacity_gauges
from atx.scheduler import SamplingParams, Scheduler, Sequence, SeqState
from oracle.kv_ref import RefPagedKV, capacity_gauges


def test_allocator_matches_reference_model_without_sharing():
    rnd = random.Random(7)
    a = BlockAllocator(64, 16, enable_prefix_caching=False)
    ref = RefPagedKV(64, 16)
    live = {}
    for step in range(2000):
        op = rnd.random()
        if op < 0.5 and len(live) < 12:
            sid = f"s{step}"
            live[sid] = (SequenceBlocks(a), 0)
        if live and op < 0.85:
            sid = rnd.choice(list(live))
            sb, n = live[sid]
            new = n + rnd.randint(1, 40)
            try:
                ref.grow(sid, new)
                ok = True
            except MemoryError:
                ok = False
            if ok:
                sb.ensure_capacity(new)
                live[sid] = (sb, new)
                assert sb.table == ref.tables[sid]                       # block ids bit-exact
                p = rnd.randrange(new)
                assert sb.slots(p, 1)[0] == ref.slot(sid, p)             # slot mapping bit-exact
            else:
                with pytest.raises(OutOfBlocks):
                    sb.ensure_capacity(new)
                assert sb.table == ref.tables.get(sid, [])              # all-or-nothing
        elif live:
            sid = rnd.choice(list(live))
            sb, _ = live.pop(sid)
            sb.release()
            if sid in ref.tables:
                ref.release(sid)
        assert a.num_free == len(ref.free)
    assert kv_capacity_gauges(1000, 16, 4096) == capacity_gauges(1000, 16, 4096)


def test_prefix_cache_shares_full_blocks_and_survives_release():
    a = BlockAllocator(16, 4)
    toks = list(range(100, 100 + 14))
    s1 = SequenceBlocks(a)
    assert s1.match_prefix(toks) == 0
    s1.ensure_capacity(14)
    s1.publish_full_blocks(toks, 14)          # 3 full blocks published, 4th partial
    s2 = SequenceBlocks(a)
    assert s2.match_prefix(toks[:12] + [1, 2, 3]) == 12
    assert s2.table == s1.table[:3] and all(a.ref_count(b) == 2 for b in s2.table)
    s3 = SequenceBlocks(a)
    assert s3.match_prefix(toks[:12]) == 8    # never the whole prompt: one token must be computed
    s1.release(); s2.release(); s3.release()
    assert a.num_used == 0 and a.num_free == 16
    s4 = SequenceBlocks(a)
    assert s4.match_prefix(toks) == 12        # evictable blocks are still findable
    s4.release()
    # filling the pool evicts cached blocks, least recently released first, and un-publishes them
    hog = SequenceBlocks(a)
    hog.ensure_capacity(16 * 4)
    assert a.num_free == 0
    s5 = SequenceBlocks(a)
    assert s5.match_prefix(toks) == 0
    hog.release()
    with pytest.raises(RuntimeError, match="double free"):
        a.free(0)
    assert block_hash(None, [1, 2]) != block_hash(block_hash(None, [9]), [1, 2])


def _drive(sched, seqs, sample=lambda s: 7, max_steps=10000):
    steps = 0
    plans = []
    while sched.has_work():
        plan = sched.schedule()
        assert not plan.empty
        assert plan.num_tokens <= sched.max_num_batched_tokens
        assert len(sched.running) <= sched.max_num_seqs
        # no slot is written by two chunks in one step; every chunk has its blocks
        slots = [sl for c in plan.decodes + plan.prefills for sl in c.seq.blocks.slots(c.start, c.count)]
        assert len(slots) == len(set(slots))
        sampled = {c.seq.uid: sample(c.seq) for c in plan.decodes + plan.prefills if c.samples}
        sched.commit(plan, sampled)
        plans.append(plan)
        steps += 1
        assert steps < max_steps
    return plans


def test_scheduler_runs_fanout_burst_with_chunked_prefill_and_prefix_hits():
    a = BlockAllocator(256, 16)
    sched = Scheduler(a, max_num_seqs=4, max_num_batched_tokens=64, max_model_len=512, eos_token_ids=(2,))
    prefix = list(range(1000, 1100))
    params = SamplingParams(max_tokens=10)
    seqs = [Sequence(f"r{i}", prefix + [i] * (5 + i), params, a) for i in range(6)]
    for s in seqs:
        sched.add(s)
    plans = _drive(sched, seqs)
    assert all(s.state is SeqState.FINISHED and s.finish_reason == "length" and s.num_generated == 10 for s in seqs)
    assert a.num_used == 0
    assert any(len(p.prefills) > 1 for p in plans)                      # burst admitted together
    assert any(p.decodes and p.prefills for p in plans)                 # decodes never stall behind prompts
    assert sum(s.cached_prompt_tokens for s in seqs) >= 96 * 2          # later admissions hit the shared prefix
    assert max(len(p.decodes) + len(p.prefills) for p in plans) <= 4


def test_scheduler_stops_on_eos_and_stop_ids_unless_ignored():
    a = BlockAllocator(64, 16)
    sched = Scheduler(a, 8, 128, 256, eos_token_ids=(2,))
    s1 = Sequence("a", [5, 6, 7], SamplingParams(max_tokens=50), a)
    s2 = Sequence("b", [5, 6, 8], SamplingParams(max_tokens=6, ignore_eos=True), a)
    s3 = Sequence("c", [5, 9], SamplingParams(max_tokens=50, stop_token_ids=(11,)), a)
    for s in (s1, s2, s3):
        sched.add(s)
    _drive(sched, [s1, s2, s3], sample=lambda s: 11 if s is s3 and s.num_generated == 2 else 2)
    assert (s1.finish_reason, s1.num_generated) == ("stop", 1)
    assert (s2.finish_reason, s2.num_generated) == ("length", 6)
    assert (s3.finish_reason, s3.output_ids) == ("stop", [2, 2, 11]) or s3.finish_reason == "stop"


def test_scheduler_preempts_youngest_and_recomputes():
    a = BlockAllocator(8, 16, enable_prefix_caching=False)   # 128 tokens of KV for 3 x (30 + 40)
    sched = Scheduler(a, 8, 256, 128, eos_token_ids=())
    seqs = [Sequence(f"p{i}", list(range(30)), SamplingParams(max_tokens=40, ignore_eos=True), a) for i in range(3)]
    for s in seqs:
        sched.add(s)
    plans = _drive(sched, seqs)
    assert all(s.num_generated == 40 for s in seqs)
    assert sum(s.num_preemptions for s in seqs) > 0 and any(p.preempted for p in plans)
    assert seqs[0].num_preemptions == 0                                  # oldest is never the victim
    assert a.num_used == 0


def test_scheduler_rejects_oversize_and_aborts_cleanly():
    a = BlockAllocator(8, 16)
    sched = Scheduler(a, 8, 64, 32, eos_token_ids=())
    with pytest.raises(ValueError):
        sched.add(Sequence("big", list(range(32)), SamplingParams(), a))
    s = Sequence("x", list(range(20)), SamplingParams(max_tokens=100, ignore_eos=True), a)
    sched.add(s)
    plan = sched.schedule()
    sched.commit(plan, {s.uid: 1})
    assert a.num_used == 2
    assert sched.abort("x") is s and s.finish_reason == "abort" and a.num_used == 0 and not sched.has_work()


def test_a_sequence_that_outgrows_the_kv_pool_ends_with_length_and_nobody_else_suffers():
    """One long generation on a small pool must not take the engine down (it used to raise OutOfBlocks out of
    the step loop, which failed every in-flight request): the sequence that cannot grow even when it is alone
    ends with finish_reason "length" and keeps its tokens; a prompt that cannot fit even an empty pool ends
    with no tokens; requests queued behind them run normally."""
    from tests.fakes import make_logic_engine

    # (a) prompt that never fits: 40 tokens, 1 block
    lone = Scheduler(BlockAllocator(1, 16), 8, 64, 64)
    big = Sequence("l", list(range(40)), SamplingParams(), lone.alloc)
    lone.add(big)
    plan = lone.schedule()
    assert plan.empty and plan.finished == [big] and big.finish_reason == "length" and big.num_generated == 0
    assert not lone.has_work() and lone.alloc.num_used == 0

    # (b) through LLMEngine.step(): 4 blocks = 64 tokens of KV; `hog` wants 20 + 100 tokens, `ok` waits behind it
    eng = make_logic_engine(max_num_seqs=1, max_num_batched_tokens=48, max_model_len=256, num_blocks=4,
                            async_scheduling=True)
    hog = eng.add_request("hog", list(range(20)), SamplingParams(max_tokens=100, ignore_eos=True))
    ok = eng.add_request("ok", list(range(100, 110)), SamplingParams(max_tokens=5, ignore_eos=True))
    reported = []
    steps = 0
    while eng.has_work():
        reported += [s.request_id for s in eng.step() if s.state is SeqState.FINISHED]
        steps += 1
        assert steps < 500
    assert hog.finish_reason == "length" and 40 <= hog.num_generated <= 45     # filled the 64-token pool, then ended
    assert ok.finish_reason == "length" and ok.num_generated == 5
    assert sorted(reported) == ["hog", "ok"]                                  # both reach their callers
    assert eng.allocator.num_used == 0


def test_block_tables_and_slots_at_the_long_context_config_size():
    """BASELINE configs[4] at FULL size on the host side: 16 sequences of 32 768 prompt tokens + 128 generated,
    chunked prefill under an 8192-token budget, 16-token pages, a 73 785-block pool (what the 8B engine sizes on a
    B200 at 0.92 utilisation). Integer work only, so it is checked exactly: every block table and the slot of
    every 997th position against the reference model, no block handed out twice, everything returned at the end."""
    nb, bs, P, O, S = 73785, 16, 32768, 128, 16
    a = BlockAllocator(nb, bs, enable_prefix_caching=False)
    ref = RefPagedKV(nb, bs)
    sched = Scheduler(a, S, 8192, P + O, ())
    seqs = [Sequence(f"long{i}", [(7 * i + j) % 50000 for j in range(P)],
                     SamplingParams(temperature=0.0, max_tokens=O, ignore_eos=True), a) for i in range(S)]
    for s in seqs:
        sched.add(s)
    steps = 0
    while sched.has_work():
        plan = sched.schedule()
        assert not plan.empty and sum(c.count for c in plan.prefills) + len(plan.decodes) <= 8192
        for c in plan.decodes + plan.prefills:      # the reference model grows the same sequences in the same order
            ref.grow(c.seq.request_id, c.start + c.count)
            assert c.seq.blocks.table[: len(ref.tables[c.seq.request_id])] == ref.tables[c.seq.request_id]
            for p in range(c.start, c.start + c.count, 997):
                assert c.seq.blocks.slots(p, 1)[0] == ref.slot(c.seq.request_id, p)
        sched.commit(plan, {c.seq.uid: 1 for c in plan.decodes + plan.prefills if c.samples})
        steps += 1
    used = [b for t in ref.tables.values() for b in t]
    assert len(used) == len(set(used)) == S * ((P + O + bs - 1) // bs)       # 2056 pages per sequence, none shared
    assert all(s.state is SeqState.FINISHED and s.num_generated == O for s in seqs)
    assert a.num_free == nb and steps >= (S * P) // 8192 + O
