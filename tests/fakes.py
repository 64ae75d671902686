"""Test doubles (tests only — never imported by the product path)."""
from __future__ import annotations

import asyncio
import zlib
from typing import List

from atx.engine import RequestOutput


class FakeAsyncEngine:
    """Stands in for atx.engine.AsyncEngine so the HTTP contract can be tested without a GPU:
    emits `max_tokens` ids derived from the prompt, one per event-loop tick."""

    vocab_size = 128256

    class ecfg:  # noqa: N801 - mimics EngineConfig attributes
        max_num_seqs = 12
        max_num_batched_tokens = 8192
        max_model_len = 4096

    def __init__(self, fail: bool = False, delay: float = 0.0):
        self.fail = fail
        self.delay = delay
        self.seen: List[dict] = []

    def kv_cache_info(self):
        return 1000, 16

    async def generate(self, prompt_ids, params, request_id):
        self.seen.append({"ids": list(prompt_ids), "params": params, "request_id": request_id})
        if self.fail:
            raise RuntimeError("boom")
        out: List[int] = []
        n = min(params.max_tokens, 5)
        for i in range(n):
            await asyncio.sleep(self.delay)
            out.append(ord("a") + (len(prompt_ids) + i) % 26)
            yield RequestOutput(request_id, list(out), out[-1:], i == n - 1, "length" if i == n - 1 else None,
                                len(prompt_ids), 0, None)


def fake_replica_worker(rank, model, ecfg, req, resp):
    """Stands in for atx.router._engine_worker in CPU tests: echoes `max_tokens` ids tagged with the
    replica rank so tests can see where a request was routed. `model == "explode"` dies on the first
    request to exercise failure propagation."""
    resp.send(("ready", rank, (100 + rank, 16), 1024))
    while True:
        msg = req.recv()
        if msg[0] == "stop":
            return
        if msg[0] == "add":
            _, rid, ids, params = msg
            if model == "explode" or (model == "explode1" and rank == 1):
                resp.send(("fatal", rank, "boom"))
                return
            if model == "vanish1" and rank == 1:     # the process dies without a word (OOM kill, segfault)
                import os

                os._exit(3)
            if len(ids) > 500:
                resp.send(("error", rid, "ValueError('too long')"))
                continue
            n = params.max_tokens
            for i in range(n):
                toks = [1000 * rank + j for j in range(i + 1)]
                resp.send(("out", rid, toks, i == n - 1, "length" if i == n - 1 else None, len(ids), 0))
            resp.send(("stats", rank, 7 + rank, [1, 2, 2]))    # KV blocks in use + the batch size of three steps


class _LogicRunner:
    """Runner double for host-logic tests: 'samples' a deterministic token per (request, position)."""

    dev = None
    num_blocks = 64
    kernel_launches = 0

    def launch(self, plan, chained=False):
        out = {}
        for c in plan.decodes + plan.prefills:
            if c.samples:
                out[c.seq.uid] = zlib.crc32(f"{c.seq.request_id}:{c.start + c.count}".encode()) % 1000
        return out

    def wait(self, handle):
        return handle or {}

    def close(self):
        pass

    def kv_cache_info(self):
        return self.num_blocks, 16


def make_logic_engine(max_num_seqs=4, max_num_batched_tokens=48, max_model_len=256, num_blocks=64,
                      async_scheduling=True):
    """An LLMEngine with the REAL scheduler, block allocator and step()/chaining logic over a runner double:
    records the plan of every step in `.trace` so two engines can be compared call for call (CPU only)."""
    from atx.config import EngineConfig, model_config_by_name
    from atx.engine import LLMEngine
    from atx.kv_cache import BlockAllocator
    from atx.scheduler import Scheduler

    eng = LLMEngine.__new__(LLMEngine)
    eng.mcfg = model_config_by_name("tiny")
    eng.ecfg = EngineConfig(max_model_len=max_model_len, max_num_seqs=max_num_seqs,
                            max_num_batched_tokens=max_num_batched_tokens, num_gpu_blocks=num_blocks,
                            async_scheduling=async_scheduling)
    eng.runner = _LogicRunner()
    eng.runner.num_blocks = num_blocks
    eng.allocator = BlockAllocator(num_blocks, 16, True)
    eng.scheduler = Scheduler(eng.allocator, max_num_seqs, max_num_batched_tokens, max_model_len, eng.mcfg.eos_token_ids)
    eng.steps = 0
    eng.batch_size_observer = None
    eng._inflight = None
    eng.trace = []
    real_launch = eng.runner.launch

    def launch(plan, chained=False):
        eng.trace.append((chained, tuple((c.seq.request_id, c.start, c.count, c.samples) for c in plan.decodes),
                          tuple((c.seq.request_id, c.start, c.count, c.samples) for c in plan.prefills),
                          tuple(tuple(c.seq.blocks.table) for c in plan.decodes + plan.prefills)))
        return real_launch(plan, chained)
    eng.runner.launch = launch
    return eng
