"""LLMEngine (synchronous step loop) and AsyncEngine (the `AsyncLLMEngine`-shaped facade).

`AsyncEngine.generate(prompt_token_ids, sampling_params, request_id)` is an async generator
yielding RequestOutput objects as tokens arrive — the same consumption pattern as
`AsyncLLMEngine.generate` in llm/serve_llm.py:527-561 (first yielded output = time to first token).
One engine thread per GPU issues every CUDA launch; the asyncio side never blocks on the device.
"""
from __future__ import annotations

import asyncio
import threading
import time
from dataclasses import dataclass, field
from typing import AsyncIterator, Callable, Dict, List, Optional, Sequence as Seq

import torch

from .config import EngineConfig, ModelConfig, ModelSource, resolve_model
from .kv_cache import BlockAllocator
from .runner import ModelRunner
from .scheduler import SamplingParams, Scheduler, Sequence, SeqState
from .weights import ModelWeights


@dataclass
class RequestOutput:
    request_id: str
    token_ids: List[int]            # all generated ids so far
    new_token_ids: List[int]
    finished: bool
    finish_reason: Optional[str]
    prompt_len: int
    num_cached_tokens: int = 0
    ttft_s: Optional[float] = None


class LLMEngine:
    def __init__(self, source: ModelSource, ecfg: EngineConfig, weights: Optional[ModelWeights] = None):
        self.mcfg: ModelConfig = source.config
        self.ecfg = ecfg
        dev = torch.device("cuda", ecfg.device)
        torch.cuda.set_device(dev)
        if weights is None:
            if source.path is not None:
                weights = ModelWeights.from_hf_dir(self.mcfg, source.path, dev, ecfg.tp_size, ecfg.tp_rank)
            else:
                weights = ModelWeights.random(self.mcfg, source.random_seed or 0, dev, tp_size=ecfg.tp_size,
                                              tp_rank=ecfg.tp_rank)
        self.runner = ModelRunner(self.mcfg, ecfg, weights)
        self.allocator = BlockAllocator(self.runner.num_blocks, ecfg.block_size, ecfg.enable_prefix_caching)
        self.scheduler = Scheduler(self.allocator, ecfg.max_num_seqs, ecfg.max_num_batched_tokens, ecfg.max_model_len,
                                   self.mcfg.eos_token_ids)
        self.steps = 0
        self.batch_size_observer: Optional[Callable[[int], None]] = None
        self._inflight = None   # (plan, launch handle) of a step whose tokens the host has not read yet

    def add_request(self, request_id: str, prompt_ids: Seq[int], params: SamplingParams) -> Sequence:
        V = self.mcfg.vocab_size
        if any((t < 0 or t >= V) for t in prompt_ids):
            raise ValueError("prompt token id outside the vocabulary")
        budget = self.ecfg.max_model_len - len(prompt_ids)
        if budget <= 0:
            raise ValueError(f"prompt of {len(prompt_ids)} tokens does not fit max_model_len={self.ecfg.max_model_len}")
        bs = self.ecfg.block_size
        if (len(prompt_ids) + bs) // bs > self.allocator.num_blocks:
            raise ValueError(f"prompt of {len(prompt_ids)} tokens exceeds the KV cache "
                             f"({self.allocator.num_blocks} blocks of {bs})")
        if params.max_tokens > budget:
            params = SamplingParams(**{**params.__dict__, "max_tokens": budget})
        seq = Sequence(request_id, prompt_ids, params, self.allocator)
        self.scheduler.add(seq)
        return seq

    def abort(self, request_id: str) -> None:
        self.scheduler.abort(request_id)

    def has_work(self) -> bool:
        return self._inflight is not None or self.scheduler.has_work()

    def _chain(self, plan):
        """The decode step that follows `plan` if nothing can change the batch: same sequences, next
        position. Built BEFORE the host knows plan's tokens so that the GPU never waits for the host
        (scheduling, staging and the token read-back of step N overlap step N+1). Returns None whenever the
        ordinary scheduler has to run: prompts in the batch or in the queue, a sequence reaching its length
        limit, or no KV block left for the next position. A sequence that stops on EOS in step N simply has
        its row of step N+1 ignored."""
        from .scheduler import Chunk, StepPlan

        sch = self.scheduler
        if plan.prefills or not plan.decodes or sch.waiting or plan.preempted:
            return None
        if len(plan.decodes) != len(sch.running):
            return None
        need = 0
        for c in plan.decodes:
            s = c.seq
            if s.state is not SeqState.RUNNING:
                return None
            if s.num_generated + 1 >= s.params.max_tokens or len(s.tokens) + 1 >= sch.max_model_len:
                return None   # finishes on length with the in-flight token: composition changes
            need += s.blocks.blocks_needed(c.start + 2)
        if need > self.allocator.num_free:
            return None
        nxt = StepPlan()
        for c in plan.decodes:
            c.seq.blocks.ensure_capacity(c.start + 2)
            nxt.decodes.append(Chunk(c.seq, c.start + 1, 1, True))
        return nxt

    def step(self) -> List[Sequence]:
        """One scheduler + forward pass. Returns sequences that received a token."""
        if self._inflight is None:
            plan = self.scheduler.schedule()
            if plan.empty:
                return list(plan.finished)   # sequences the scheduler had to end (they outgrew the KV pool)
            if self.batch_size_observer is not None:
                self.batch_size_observer(len(plan.decodes) + len(plan.prefills))
            self._inflight = (plan, self.runner.launch(plan))
        plan, handle = self._inflight
        nxt = None
        if self.ecfg.async_scheduling:
            cplan = self._chain(plan)
            if cplan is not None:
                if self.batch_size_observer is not None:
                    self.batch_size_observer(len(cplan.decodes))
                nxt = (cplan, self.runner.launch(cplan, chained=True))
        sampled = self.runner.wait(handle)
        self.steps += 1
        touched = self.scheduler.commit(plan, sampled)
        self._inflight = nxt
        return touched + plan.finished

    def generate_sync(self, prompts: List[Seq[int]], params: SamplingParams) -> List[List[int]]:
        """Blocking helper for tests and the bench: run the given prompts to completion."""
        seqs = [self.add_request(f"sync-{i}-{id(prompts)}", p, params) for i, p in enumerate(prompts)]
        while self.has_work():
            self.step()
        return [s.output_ids for s in seqs]


class _Stream:
    __slots__ = ("queue", "loop", "seq", "submitted")

    def __init__(self, loop: asyncio.AbstractEventLoop):
        self.queue: asyncio.Queue = asyncio.Queue()
        self.loop = loop
        self.seq: Optional[Sequence] = None
        self.submitted = time.monotonic()


class AsyncEngine:
    """Thread-backed engine with an asyncio streaming interface (one per GPU / replica)."""

    def __init__(self, engine: LLMEngine, tp_group=None):
        """tp_group (atx.tp_group.TPGroup): `engine` is rank 0 of a tensor-parallel group; every batch of
        add/abort messages is sent to the follower ranks before it is applied here, one send per loop
        iteration, so all ranks make the same calls in the same order."""
        self.engine = engine
        self._tp = tp_group
        self._inbox: List = []
        self._cv = threading.Condition()
        self._streams: Dict[str, _Stream] = {}
        self._stop = False
        self._error: Optional[BaseException] = None
        self._thread = threading.Thread(target=self._run, name="atx-engine", daemon=True)
        self._thread.start()

    # ---- asyncio side -----------------------------------------------------------------------------
    async def generate(self, prompt_ids: Seq[int], params: SamplingParams, request_id: str
                       ) -> AsyncIterator[RequestOutput]:
        if self._error is not None:
            raise RuntimeError(f"engine thread died: {self._error!r}")
        loop = asyncio.get_running_loop()
        st = _Stream(loop)
        with self._cv:
            if request_id in self._streams:
                raise ValueError(f"duplicate request id {request_id!r}")
            self._streams[request_id] = st
            self._inbox.append(("add", request_id, list(prompt_ids), params))
            self._cv.notify()
        try:
            while True:
                item = await st.queue.get()
                if isinstance(item, BaseException):
                    raise item
                yield item
                if item.finished:
                    return
        finally:
            with self._cv:
                if self._streams.pop(request_id, None) is not None and (st.seq is None or
                                                                         st.seq.state is not SeqState.FINISHED):
                    self._inbox.append(("abort", request_id))  # client went away: free the KV blocks
                    self._cv.notify()

    def shutdown(self) -> None:
        with self._cv:
            self._stop = True
            self._cv.notify()
        self._thread.join(timeout=10)
        if not self._thread.is_alive():
            if self._tp is not None:
                self._tp.shutdown()
            else:
                self.engine.runner.close()

    # ---- facts the HTTP layer exports ---------------------------------------------------------------
    @property
    def ecfg(self) -> EngineConfig:
        return self.engine.ecfg

    @property
    def vocab_size(self) -> int:
        return self.engine.mcfg.vocab_size

    def kv_cache_info(self):
        return self.engine.runner.kv_cache_info()

    def set_batch_size_observer(self, fn) -> None:
        self.engine.batch_size_observer = fn

    def stats(self) -> dict:
        a = self.engine.allocator
        return {"kv_used_blocks": a.num_used, "prefix_hits": a.hits, "prefix_queries": a.queries,
                "steps": self.engine.steps, "kernel_launches": self.engine.runner.kernel_launches}

    # ---- engine thread ----------------------------------------------------------------------------
    def _post(self, st: _Stream, item) -> None:
        st.loop.call_soon_threadsafe(st.queue.put_nowait, item)

    def _run(self) -> None:
        eng = self.engine
        try:
            if eng.runner.dev is not None:
                torch.cuda.set_device(eng.runner.dev)
            while True:
                with self._cv:
                    while not self._inbox and not eng.has_work() and not self._stop:
                        self._cv.wait()
                    if self._stop:
                        return
                    inbox, self._inbox = self._inbox, []
                if self._tp is not None:
                    self._tp.broadcast(inbox)   # followers apply the same messages, then step iff they have work
                for msg in inbox:
                    if msg[0] == "add":
                        _, rid, ids, params = msg
                        st = self._streams.get(rid)
                        try:   # a request whose client already left is added all the same (its abort follows):
                            seq = eng.add_request(rid, ids, params)   # every rank must see the same calls
                            if st is not None:
                                st.seq = seq
                        except Exception as exc:  # bad request: surface to the awaiting handler
                            if st is not None:
                                self._post(st, exc)
                    else:
                        eng.abort(msg[1])
                if not eng.has_work():
                    continue
                for seq in eng.step():
                    st = self._streams.get(seq.request_id)
                    if st is None:
                        continue
                    fin = seq.state is SeqState.FINISHED
                    n_out = seq.num_generated
                    # every cross-thread wake-up costs the engine thread the GIL: non-streaming callers only
                    # need the first token (TTFT), an occasional progress mark and the final result
                    if not (fin or n_out == 1 or seq.params.stream or n_out % 32 == 0):
                        continue
                    out = seq.output_ids
                    ttft = (seq.first_token_time - seq.arrival) if seq.first_token_time else None
                    self._post(st, RequestOutput(seq.request_id, out, out[-1:], fin, seq.finish_reason,
                                                 seq.prompt_len, seq.cached_prompt_tokens, ttft))
        except BaseException as exc:  # device fault etc.: fail every waiter loudly, never hang them
            self._error = exc
            with self._cv:
                streams = list(self._streams.values())
            for st in streams:
                self._post(st, RuntimeError(f"engine failure: {exc!r}"))


def build_engine(model: str, ecfg: EngineConfig, weights: Optional[ModelWeights] = None) -> LLMEngine:
    return LLMEngine(resolve_model(model), ecfg, weights)
