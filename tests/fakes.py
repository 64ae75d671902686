"""Test doubles (tests only — never imported by the product path)."""
from __future__ import annotations

import asyncio
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
            if model == "explode":
                resp.send(("fatal", rank, "boom"))
                return
            if len(ids) > 500:
                resp.send(("error", rid, "ValueError('too long')"))
                continue
            n = params.max_tokens
            for i in range(n):
                toks = [1000 * rank + j for j in range(i + 1)]
                resp.send(("out", rid, toks, i == n - 1, "length" if i == n - 1 else None, len(ids), 0))
