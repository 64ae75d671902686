"""Trivially-correct model of paged-KV bookkeeping (TEST INFRASTRUCTURE — see llama_ref.py header).

The reference delegates this to vLLM's block manager (`AsyncLLMEngine`, llm/serve_llm.py:378) and
only observes its capacity (`num_gpu_blocks`, `block_size`; llm/serve_llm.py:209-221, 410-422), so
there is nothing to pin against beyond those formulas; this file states the invariants our
allocator must satisfy and a dict-based model whose block ids are compared bit-exactly in the
no-sharing case (FIFO free list)."""
from __future__ import annotations

from collections import deque
from typing import Dict, List


class RefPagedKV:
    """No prefix sharing: blocks handed out FIFO from a free queue, returned on release in reverse
    table order (so the tail block is reused first only after the queue wraps)."""

    def __init__(self, num_blocks: int, block_size: int):
        self.free = deque(range(num_blocks))
        self.bs = block_size
        self.tables: Dict[str, List[int]] = {}

    def grow(self, sid: str, new_len: int) -> None:
        t = self.tables.setdefault(sid, [])
        need = (new_len + self.bs - 1) // self.bs - len(t)
        if need > len(self.free):
            raise MemoryError
        for _ in range(max(0, need)):
            t.append(self.free.popleft())

    def release(self, sid: str) -> None:
        for b in reversed(self.tables.pop(sid)):
            self.free.append(b)

    def slot(self, sid: str, pos: int) -> int:
        return self.tables[sid][pos // self.bs] * self.bs + pos % self.bs


def capacity_gauges(num_blocks: int, block_size: int, max_model_len: int):
    """llm/serve_llm.py:215-221 — total tokens, floor estimate, exact ratio."""
    total = num_blocks * block_size
    return total, total // max_model_len, total / max_model_len
