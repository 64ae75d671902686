"""Continuous-batching scheduler with chunked prefill, prefix-cache admission and preemption.

Replaces the vLLM V1 scheduler the reference reaches through `AsyncLLMEngine.generate`
(llm/serve_llm.py:527) under the two caps it configures: `max_num_seqs` and
`max_num_batched_tokens` (llm/serve_llm.py:367-370). Pure host logic, no torch.

One step = one forward pass over a flat token batch:
  * every RUNNING sequence that has finished its prompt contributes 1 decode token,
  * the remaining token budget is filled with prompt chunks of running / newly admitted
    sequences (FCFS), so a burst of agent requests that arrives together is prefetched together
    and long prompts never stall the decodes already in flight,
  * when the KV pool cannot grow a running sequence, the most recently admitted sequence is
    preempted (blocks released, recomputed later — its full blocks usually survive in the prefix
    cache).
"""
from __future__ import annotations

import enum
import hashlib
import itertools
import time
from collections import deque
from dataclasses import dataclass, field
from typing import Deque, Dict, List, Optional, Sequence as Seq, Tuple

from .kv_cache import BlockAllocator, OutOfBlocks, SequenceBlocks


@dataclass
class SamplingParams:
    """vLLM-compatible subset. The reference uses temperature=0.2, max_tokens (llm/serve_llm.py:521)."""

    temperature: float = 0.2
    top_p: float = 1.0
    max_tokens: int = 512
    seed: Optional[int] = None
    ignore_eos: bool = False
    stop_token_ids: Tuple[int, ...] = ()
    stream: bool = False   # deliver every token to the caller (else: first token, every 32nd, and the last)


def default_seed(request_id: str) -> int:
    """Sampling seed of a request that did not bring one: a hash of its id that is the SAME in every process
    (Python's own str hash is salted per process). Tensor-parallel ranks build their Sequence objects
    independently from the broadcast request stream and must draw identical noise from identical logits, or
    their KV caches and residual shards diverge; it also makes a restarted server repeat itself."""
    digest = hashlib.blake2b(request_id.encode("utf-8", "surrogatepass"), digest_size=8).digest()
    return int.from_bytes(digest, "little") & 0x7FFFFFFFFFFFFFFF


class SeqState(enum.Enum):
    WAITING = 0
    RUNNING = 1
    FINISHED = 2


_seq_counter = itertools.count()


class Sequence:
    def __init__(self, request_id: str, prompt_ids: Seq[int], params: SamplingParams, allocator: BlockAllocator,
                 arrival: Optional[float] = None):
        if len(prompt_ids) == 0:
            raise ValueError("empty prompt")
        self.request_id = request_id
        self.uid = next(_seq_counter)
        self.prompt_len = len(prompt_ids)
        self.tokens: List[int] = list(prompt_ids)   # prompt + generated
        self.params = params
        self.blocks = SequenceBlocks(allocator)
        self.num_computed = 0                        # tokens whose K/V are in the cache
        self.state = SeqState.WAITING
        self.finish_reason: Optional[str] = None
        self.arrival = time.monotonic() if arrival is None else arrival
        self.first_token_time: Optional[float] = None
        self.seed = params.seed if params.seed is not None else default_seed(request_id)
        self.num_preemptions = 0
        self.cached_prompt_tokens = 0               # prompt tokens served by the prefix cache

    @property
    def num_generated(self) -> int:
        return len(self.tokens) - self.prompt_len

    @property
    def output_ids(self) -> List[int]:
        return self.tokens[self.prompt_len:]

    @property
    def in_prefill(self) -> bool:
        """True while some already-known token still lacks K/V other than the newest one."""
        return self.num_computed < len(self.tokens) - 1 or self.num_generated == 0


@dataclass
class Chunk:
    seq: Sequence
    start: int          # first token position processed this step
    count: int          # number of tokens processed
    samples: bool       # the last token of the chunk produces the next token


@dataclass
class StepPlan:
    decodes: List[Chunk] = field(default_factory=list)
    prefills: List[Chunk] = field(default_factory=list)
    preempted: List[Sequence] = field(default_factory=list)
    finished: List[Sequence] = field(default_factory=list)   # ended by the scheduler itself (KV pool too small)

    @property
    def num_tokens(self) -> int:
        return sum(c.count for c in self.decodes) + sum(c.count for c in self.prefills)

    @property
    def empty(self) -> bool:
        return not self.decodes and not self.prefills


class Scheduler:
    def __init__(self, allocator: BlockAllocator, max_num_seqs: int, max_num_batched_tokens: int, max_model_len: int,
                 eos_token_ids: Seq[int] = ()):
        self.alloc = allocator
        self.max_num_seqs = max_num_seqs
        self.max_num_batched_tokens = max_num_batched_tokens
        self.max_model_len = max_model_len
        self.eos = set(eos_token_ids)
        self.waiting: Deque[Sequence] = deque()
        self.running: List[Sequence] = []

    # ---- queue management ---------------------------------------------------------------------------
    def add(self, seq: Sequence) -> None:
        if seq.prompt_len >= self.max_model_len:
            raise ValueError(f"prompt of {seq.prompt_len} tokens does not fit max_model_len={self.max_model_len}")
        self.waiting.append(seq)

    def abort(self, request_id: str) -> Optional[Sequence]:
        for q in (self.running, self.waiting):
            for s in list(q):
                if s.request_id == request_id:
                    q.remove(s)
                    self._finish(s, "abort")
                    return s
        return None

    def has_work(self) -> bool:
        return bool(self.waiting or self.running)

    def _finish(self, seq: Sequence, reason: str) -> None:
        seq.state = SeqState.FINISHED
        seq.finish_reason = reason
        seq.blocks.release()

    def _preempt(self, seq: Sequence, plan: StepPlan) -> None:
        """Recompute-style preemption: drop the blocks, keep the tokens, go back to the queue head."""
        self.running.remove(seq)
        seq.blocks.release()
        seq.num_computed = 0
        seq.state = SeqState.WAITING
        seq.num_preemptions += 1
        self.waiting.appendleft(seq)
        plan.preempted.append(seq)

    # ---- one step -------------------------------------------------------------------------------------
    def schedule(self) -> StepPlan:
        plan = StepPlan()
        budget = self.max_num_batched_tokens

        # 1) decodes of running sequences (oldest first); preempt the youngest on KV exhaustion
        for seq in list(self.running):
            if seq.state is not SeqState.RUNNING or seq.in_prefill:
                continue
            if budget <= 0:
                break
            while True:
                try:
                    seq.blocks.ensure_capacity(len(seq.tokens))
                    break
                except OutOfBlocks:
                    victim = self.running[-1]
                    if victim is seq and len(self.running) == 1:
                        # alone in the pool and still out of blocks: this sequence has outgrown the KV cache.
                        # It ends here with what it has ("length"), the engine and every other request live on.
                        self._retire_oversized(seq, plan)
                        break
                    self._preempt(victim, plan)
                    plan.decodes = [c for c in plan.decodes if c.seq is not victim]
                    if victim is seq:
                        break
            if seq.state is not SeqState.RUNNING:
                continue
            plan.decodes.append(Chunk(seq, len(seq.tokens) - 1, 1, True))
            budget -= 1

        # 2) prompt chunks: running sequences still in prefill, then admissions from the queue
        def add_prefill(seq: Sequence, budget: int) -> int:
            remaining = len(seq.tokens) - seq.num_computed
            count = min(remaining, budget)
            if count <= 0:
                return 0
            try:
                seq.blocks.ensure_capacity(seq.num_computed + count)
            except OutOfBlocks:
                return -1
            plan.prefills.append(Chunk(seq, seq.num_computed, count, count == remaining))
            # Publish the blocks this chunk will fill BEFORE the forward pass: requests admitted later
            # in this same step (a fan-out burst arrives together) attach to them instead of
            # recomputing the shared prefix. Safe because the K/V writes of every token in the batch
            # precede the attention reads of that layer, and nothing is preempted after this point.
            seq.blocks.publish_full_blocks(seq.tokens, seq.num_computed + count)
            return count

        for seq in self.running:
            if budget <= 0:
                break
            if seq.state is SeqState.RUNNING and seq.in_prefill and not any(c.seq is seq for c in plan.decodes):
                used = add_prefill(seq, budget)
                if used > 0:
                    budget -= used

        while self.waiting and budget > 0 and len(self.running) < self.max_num_seqs and not plan.preempted:
            seq = self.waiting[0]
            if not seq.blocks.table:
                cached = seq.blocks.match_prefix(seq.tokens)
                seq.num_computed = cached
                seq.cached_prompt_tokens = cached
            used = add_prefill(seq, budget)
            if used <= 0:
                if used < 0:  # no room in the KV pool right now: give the matched prefix back and wait
                    seq.blocks.release()
                    seq.num_computed = 0
                break
            self.waiting.popleft()
            seq.state = SeqState.RUNNING
            self.running.append(seq)
            budget -= used
        if plan.empty and not self.running and self.waiting:
            # nothing else holds a block and the head of the queue still does not fit: it never will
            self._retire_oversized(self.waiting.popleft(), plan)
            return self._again(plan)
        if plan.empty and self.running:
            # every running sequence is stuck mid-prompt on an exhausted pool: evict the youngest
            if len(self.running) == 1:
                self._retire_oversized(self.running[0], plan)
            else:
                self._preempt(self.running[-1], plan)
            return self._again(plan)
        return plan

    def _retire_oversized(self, seq: Sequence, plan: StepPlan) -> None:
        if seq in self.running:
            self.running.remove(seq)
        self._finish(seq, "length")
        plan.finished.append(seq)

    def _again(self, plan: StepPlan) -> StepPlan:
        if not self.has_work():
            return plan
        again = self.schedule()
        again.preempted = plan.preempted + again.preempted
        again.finished = plan.finished + again.finished
        return again

    # ---- after the forward pass -----------------------------------------------------------------------
    def commit(self, plan: StepPlan, sampled: Dict[int, int], now: Optional[float] = None) -> List[Sequence]:
        """Record computed tokens, append sampled ids (keyed by Sequence.uid), retire finished
        sequences. Returns the sequences that received a token this step."""
        now = time.monotonic() if now is None else now
        touched: List[Sequence] = []
        for c in plan.decodes + plan.prefills:
            seq = c.seq
            if seq.state is not SeqState.RUNNING:
                continue
            seq.num_computed = c.start + c.count
            seq.blocks.publish_full_blocks(seq.tokens, seq.num_computed)
            if not c.samples:
                continue
            tok = int(sampled[seq.uid])
            seq.tokens.append(tok)
            if seq.first_token_time is None:
                seq.first_token_time = now
            touched.append(seq)
            p = seq.params
            if not p.ignore_eos and (tok in self.eos or tok in p.stop_token_ids):
                reason = "stop"
            elif seq.num_generated >= p.max_tokens:
                reason = "length"
            elif len(seq.tokens) >= self.max_model_len:
                reason = "length"
            else:
                reason = None
            if reason is not None:
                self.running.remove(seq)
                self._finish(seq, reason)
        return touched
