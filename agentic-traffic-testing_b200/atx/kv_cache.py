"""Paged KV-cache block allocator with reference counts and a block-hash prefix cache.

Host-side replacement for vLLM's KV block manager that sits under the reference's
`AsyncLLMEngine` (llm/serve_llm.py:378; capacity is what `_kv_cache_info` :410 and
`_update_kv_cache_gauges` :209-221 export). All of it is integer bookkeeping and is checked
bit-exactly against oracle/kv_ref.py.

Layout on the device (owned by ModelRunner): K and V pools of shape
[num_blocks, n_kv_heads, block_size, head_dim] bf16 per layer; token position p of a sequence with
block table `bt` lives in slot `bt[p // block_size] * block_size + p % block_size`.

Prefix cache: a FULL block is identified by the hash chain
key_i = H(key_{i-1}, tokens of block i); sequences whose prompts share a prefix (the agent fan-out
shares template + task + decision text) map those blocks to the same physical block
(ref-counted, copy never needed because full blocks are immutable). Blocks whose ref-count drops
to zero stay findable in an LRU list until the allocator needs them back.
"""
from __future__ import annotations

import hashlib
import struct
from collections import OrderedDict, deque
from typing import Deque, Dict, List, Optional, Sequence, Tuple


class OutOfBlocks(RuntimeError):
    pass


def block_hash(parent: Optional[bytes], tokens: Sequence[int]) -> bytes:
    """Content hash of one full block given its parent's hash (collision-safe: 128-bit blake2b)."""
    h = hashlib.blake2b(digest_size=16)
    h.update(parent or b"\x00" * 16)
    h.update(struct.pack(f"<{len(tokens)}i", *tokens))
    return h.digest()


class BlockAllocator:
    def __init__(self, num_blocks: int, block_size: int, enable_prefix_caching: bool = True):
        if num_blocks <= 0 or block_size <= 0:
            raise ValueError("num_blocks and block_size must be positive")
        self.num_blocks = num_blocks
        self.block_size = block_size
        self.enable_prefix_caching = enable_prefix_caching
        self._free: Deque[int] = deque(range(num_blocks))          # never-hashed / reclaimed blocks
        self._ref: List[int] = [0] * num_blocks
        self._hash_of: List[Optional[bytes]] = [None] * num_blocks  # hash of a cached full block
        self._cached: Dict[bytes, int] = {}                          # hash -> block id
        self._evictable: "OrderedDict[int, None]" = OrderedDict()    # cached blocks with ref 0, LRU order
        self.hits = 0
        self.queries = 0

    # ---- capacity -----------------------------------------------------------------------------
    @property
    def num_free(self) -> int:
        return len(self._free) + len(self._evictable)

    @property
    def num_used(self) -> int:
        return self.num_blocks - self.num_free

    def can_allocate(self, n: int) -> bool:
        return n <= self.num_free

    # ---- primitives ---------------------------------------------------------------------------
    def allocate(self) -> int:
        """A private (un-hashed) block with ref-count 1."""
        if self._free:
            b = self._free.popleft()
        elif self._evictable:
            b, _ = self._evictable.popitem(last=False)  # least recently released cached block
            del self._cached[self._hash_of[b]]          # type: ignore[arg-type]
            self._hash_of[b] = None
        else:
            raise OutOfBlocks("KV cache exhausted")
        self._ref[b] = 1
        return b

    def free(self, block: int) -> None:
        if self._ref[block] <= 0:
            raise RuntimeError(f"double free of KV block {block}")
        self._ref[block] -= 1
        if self._ref[block] == 0:
            if self._hash_of[block] is not None:
                self._evictable[block] = None  # keep findable
            else:
                self._free.append(block)

    def lookup(self, key: bytes) -> Optional[int]:
        """Take a reference on the cached block with this hash, if any."""
        if not self.enable_prefix_caching:
            return None
        self.queries += 1
        b = self._cached.get(key)
        if b is None:
            return None
        if self._ref[b] == 0:
            self._evictable.pop(b, None)
        self._ref[b] += 1
        self.hits += 1
        return b

    def publish(self, block: int, key: bytes) -> int:
        """Mark a now-full private block as cached under `key`.

        If another block already holds this content the caller keeps using its own (no copy, no
        remap of in-flight kernels); only the first publisher becomes findable."""
        if not self.enable_prefix_caching or self._hash_of[block] is not None:
            return block
        if key not in self._cached:
            self._cached[key] = block
            self._hash_of[block] = key
        return block

    def ref_count(self, block: int) -> int:
        return self._ref[block]

    def reset_prefix_cache(self) -> None:
        for b in list(self._evictable):
            self._hash_of[b] = None
            self._free.append(b)
        self._evictable.clear()
        for key, b in list(self._cached.items()):
            if self._ref[b] == 0:
                del self._cached[key]


class SequenceBlocks:
    """Block table of one sequence plus the hash chain of its full blocks."""

    def __init__(self, allocator: BlockAllocator):
        self.alloc = allocator
        self.table: List[int] = []
        self.hashes: List[bytes] = []   # one per FULL, published-or-matched block
        self.num_cached_tokens = 0      # prefix served from the cache at admission

    def match_prefix(self, tokens: Sequence[int]) -> int:
        """Attach cached blocks for the longest cached prefix of `tokens` (full blocks only, and
        never the whole prompt: at least one token must be computed to produce logits)."""
        assert not self.table
        bs = self.alloc.block_size
        max_full = (len(tokens) - 1) // bs
        parent: Optional[bytes] = None
        for i in range(max_full):
            key = block_hash(parent, tokens[i * bs:(i + 1) * bs])
            b = self.alloc.lookup(key)
            if b is None:
                break
            self.table.append(b)
            self.hashes.append(key)
            parent = key
        self.num_cached_tokens = len(self.table) * bs
        return self.num_cached_tokens

    def blocks_needed(self, new_len: int) -> int:
        bs = self.alloc.block_size
        return max(0, (new_len + bs - 1) // bs - len(self.table))

    def ensure_capacity(self, new_len: int) -> None:
        """Grow the table so that positions [0, new_len) have slots (all-or-nothing)."""
        need = self.blocks_needed(new_len)
        if not self.alloc.can_allocate(need):
            raise OutOfBlocks(f"need {need} blocks, {self.alloc.num_free} free")
        for _ in range(need):
            self.table.append(self.alloc.allocate())

    def publish_full_blocks(self, tokens: Sequence[int], num_computed: int) -> None:
        """After `num_computed` tokens have their K/V written, make newly full blocks findable."""
        bs = self.alloc.block_size
        full = num_computed // bs
        while len(self.hashes) < full:
            i = len(self.hashes)
            key = block_hash(self.hashes[-1] if self.hashes else None, tokens[i * bs:(i + 1) * bs])
            self.alloc.publish(self.table[i], key)
            self.hashes.append(key)

    def slots(self, start: int, count: int) -> List[int]:
        bs = self.alloc.block_size
        t = self.table
        return [t[p // bs] * bs + p % bs for p in range(start, start + count)]

    def release(self) -> None:
        for b in reversed(self.table):  # tail blocks become evictable before the shared prefix
            self.alloc.free(b)
        self.table = []
        self.hashes = []
        self.num_cached_tokens = 0


def kv_capacity_gauges(num_blocks: int, block_size: int, max_model_len: int) -> Tuple[int, int, float]:
    """(total_tokens, est_max_concurrency, computed_max_concurrency) — llm/serve_llm.py:215-221."""
    total = num_blocks * block_size
    return total, (total // max_model_len if max_model_len > 0 else 0), (
        total / max_model_len if max_model_len > 0 else 0.0)
