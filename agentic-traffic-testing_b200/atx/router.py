"""Data-parallel replicas behind one front door ("replicas only").

The reference is single-GPU (`count: 1`, infra/docker-compose.yml:46-52); BASELINE.json asks for the
8B model as one replica per GPU on 2/4/8 B200s. Requests are independent, so there is NO collective on
the data path: one engine *process* per GPU (own weights, own KV pool, own CUDA context and GIL), a
pair of pipes per replica, and a router in the HTTP process:

  * prefix affinity — the hash of a prompt's first blocks remembers which replica served it last, so a
    fan-out burst that shares template + task text lands on the replica whose prefix cache is hot;
  * least outstanding tokens — otherwise, and whenever the preferred replica is clearly busier.

`ReplicaPool` exposes the same `generate()` async iterator as `AsyncEngine`, so `B200Backend` does not
care which one it holds (LLM_DP_SIZE > 1 selects the pool).
"""
from __future__ import annotations

import asyncio
import multiprocessing as mp
import queue
import threading
from typing import AsyncIterator, Callable, Dict, List, Optional, Sequence as Seq

from .config import EngineConfig
from .kv_cache import block_hash
from .scheduler import SamplingParams

AFFINITY_BLOCKS = 4          # prompt blocks hashed for affinity (64 tokens at block_size 16)
SPILL_FACTOR = 2.0           # leave the preferred replica when it holds this many times the lightest load ...
SPILL_SLACK = 8192           # ... plus this many outstanding tokens: one fan-out burst (6 x ~1.3k) stays together even
#                              when other replicas are idle (its requests share a prefix and decode in one batch at
#                              the cost of one), the NEXT burst goes to the lightest replica


def _engine_worker(rank: int, model: str, ecfg: EngineConfig, req, resp) -> None:
    """Replica process: an LLMEngine on cuda:<rank> driven by messages from the front door."""
    import dataclasses

    import torch

    from .engine import build_engine
    from .scheduler import SeqState

    try:
        eng = build_engine(model, dataclasses.replace(ecfg, device=rank))
        resp.send(("ready", rank, eng.runner.kv_cache_info(), eng.mcfg.vocab_size))
    except BaseException as exc:  # no GPU / no library: report, never hang the parent
        resp.send(("fatal", rank, repr(exc)))
        return
    batch_sizes: List[int] = []          # scheduler batch size of every step since the last stats message
    eng.batch_size_observer = batch_sizes.append
    while True:
        block = not eng.has_work()
        while req.poll(None if block else 0):
            msg = req.recv()
            block = False
            if msg[0] == "stop":
                return
            if msg[0] == "add":
                try:
                    eng.add_request(msg[1], msg[2], msg[3])
                except Exception as exc:
                    resp.send(("error", msg[1], repr(exc)))
            elif msg[0] == "abort":
                eng.abort(msg[1])
        if not eng.has_work():
            continue
        try:
            for seq in eng.step():
                fin = seq.state is SeqState.FINISHED
                if not (fin or seq.num_generated == 1 or seq.params.stream or seq.num_generated % 32 == 0):
                    continue
                out = seq.output_ids
                resp.send(("out", seq.request_id, out, fin, seq.finish_reason, seq.prompt_len,
                           seq.cached_prompt_tokens))
            if eng.steps % 16 == 0 or not eng.has_work():
                resp.send(("stats", rank, eng.allocator.num_used, batch_sizes[:]))
                batch_sizes.clear()
        except BaseException as exc:
            resp.send(("fatal", rank, repr(exc)))
            return


class Router:
    """Pure routing policy (unit-testable without processes)."""

    def __init__(self, n: int, block_size: int):
        self.n, self.bs = n, block_size
        self.load = [0] * n                      # outstanding tokens (prompt + max new) per replica
        self.alive = [True] * n                  # a replica whose process died takes no more requests
        self._home: Dict[bytes, int] = {}

    def _key(self, ids: Seq[int]) -> Optional[bytes]:
        k = min(AFFINITY_BLOCKS, len(ids) // self.bs)
        if k == 0:
            return None
        key = None
        for i in range(k):
            key = block_hash(key, ids[i * self.bs:(i + 1) * self.bs])
        return key

    def pick(self, ids: Seq[int], cost: int) -> int:
        live = [r for r in range(self.n) if self.alive[r]]
        if not live:
            raise RuntimeError("no engine replica is alive")
        lightest = min(live, key=lambda r: (self.load[r], r))
        key = self._key(ids)
        choice = lightest
        if key is not None:
            home = self._home.get(key)
            if home is not None and self.alive[home] and self.load[home] <= SPILL_FACTOR * self.load[lightest] + max(cost, SPILL_SLACK):
                choice = home
            self._home[key] = choice
            if len(self._home) > 65536:
                self._home.pop(next(iter(self._home)))
        self.load[choice] += cost
        return choice

    def done(self, replica: int, cost: int) -> None:
        self.load[replica] = max(0, self.load[replica] - cost)

    def mark_dead(self, replica: int) -> None:
        self.alive[replica] = False
        self.load[replica] = 0


class ReplicaPool:
    def __init__(self, model: str, ecfg: EngineConfig, n: int, worker: Callable = _engine_worker):
        from .engine import RequestOutput

        self._RequestOutput = RequestOutput
        self.ecfg = ecfg
        self.n = n
        ctx = mp.get_context("spawn")
        self._req, self._resp, self._procs = [], [], []
        for r in range(n):
            a_recv, a_send = ctx.Pipe(duplex=False)   # front door -> replica
            b_recv, b_send = ctx.Pipe(duplex=False)   # replica -> front door
            p = ctx.Process(target=worker, args=(r, model, ecfg, a_recv, b_send), daemon=True, name=f"atx-replica-{r}")
            p.start()
            self._req.append(a_send)
            self._resp.append(b_recv)
            self._procs.append(p)
        self._kv = [(0, ecfg.block_size)] * n
        self.vocab_size = 128256
        for r in range(n):  # wait for every replica to finish loading (fail loudly if one cannot)
            msg = self._resp[r].recv()
            if msg[0] != "ready":
                self.shutdown()
                raise RuntimeError(f"replica {r} failed to start: {msg[2]}")
            self._kv[r], self.vocab_size = msg[2], msg[3]
        self.router = Router(n, ecfg.block_size)
        self._streams: Dict[str, tuple] = {}          # rid -> (loop, queue, replica, cost, token list)
        self._lock = threading.Lock()
        self._used = [0] * n
        self._batch_observer = None
        self._stop = False
        # pipe writes happen on one writer thread per replica: a busy replica with a full pipe must never block
        # the HTTP server's event loop (the 64 KiB pipe buffer holds only a few long prompts)
        self._outq: List[queue.Queue] = [queue.Queue() for _ in range(n)]
        self._readers = [threading.Thread(target=self._read, args=(r,), daemon=True) for r in range(n)]
        self._writers = [threading.Thread(target=self._write, args=(r,), daemon=True) for r in range(n)]
        for t in self._readers + self._writers:
            t.start()

    # ---- front-door API (same as AsyncEngine) ----------------------------------------------------------
    async def generate(self, prompt_ids: Seq[int], params: SamplingParams, request_id: str) -> AsyncIterator:
        loop = asyncio.get_running_loop()
        q: asyncio.Queue = asyncio.Queue()
        cost = len(prompt_ids) + params.max_tokens
        with self._lock:
            if request_id in self._streams:
                raise ValueError(f"duplicate request id {request_id!r}")
            rep = self.router.pick(prompt_ids, cost)      # raises when every replica has died
            self._streams[request_id] = (loop, q, rep, cost, [])
        self._outq[rep].put(("add", request_id, list(prompt_ids), params))
        finished = False
        try:
            while True:
                item = await q.get()
                if isinstance(item, BaseException):
                    raise item
                yield item
                if item.finished:
                    finished = True
                    return
        finally:
            with self._lock:
                mine = self._streams.pop(request_id, None)
                if mine is not None and self.router.alive[rep]:
                    self.router.done(rep, cost)
            if mine is not None and not finished and self.router.alive[rep]:
                self._outq[rep].put(("abort", request_id))

    def _write(self, r: int) -> None:
        conn = self._req[r]
        while True:
            msg = self._outq[r].get()
            if msg is None:
                return
            try:
                conn.send(msg)
            except (OSError, BrokenPipeError, ValueError) as exc:
                self._replica_died(r, f"pipe to replica closed: {exc!r}")
                return

    def _replica_died(self, r: int, why: str) -> None:
        """Fail every request the replica holds, and route nothing more to it."""
        with self._lock:
            if not self.router.alive[r]:
                return
            self.router.mark_dead(r)
            mine = [s for s in self._streams.values() if s[2] == r]
        print(f"[atx] replica {r} is gone ({why}); {sum(self.router.alive)} of {self.n} left", flush=True)
        for loop, q, *_ in mine:
            loop.call_soon_threadsafe(q.put_nowait, RuntimeError(f"replica {r} failed: {why}"))

    def _read(self, r: int) -> None:
        conn = self._resp[r]
        while not self._stop:
            try:
                msg = conn.recv()
            except (EOFError, OSError):
                msg = ("fatal", r, "replica process exited")
            kind = msg[0]
            if kind == "out":
                _, rid, toks, fin, reason, plen, cached = msg
                with self._lock:
                    st = self._streams.get(rid)
                if st is None:
                    continue
                loop, q, _, _, acc = st
                acc[:] = toks   # replicas send the full id list (sparse updates for non-streaming callers)
                out = self._RequestOutput(rid, list(acc), list(toks[-1:]), fin, reason, plen, cached, None)
                loop.call_soon_threadsafe(q.put_nowait, out)
            elif kind == "error":
                with self._lock:
                    st = self._streams.get(msg[1])
                if st is not None:
                    st[0].call_soon_threadsafe(st[1].put_nowait, ValueError(msg[2]))
            elif kind == "stats":
                self._used[r] = msg[2]
                obs = self._batch_observer
                if obs is not None and len(msg) > 3:
                    for b in msg[3]:
                        obs(b)
            elif kind == "fatal":
                if not self._stop:
                    self._replica_died(r, msg[2])
                self._outq[r].put(None)
                return

    # ---- facts for the gauges: per-replica values summed (documented in DESIGN.md) -------------------------
    def kv_cache_info(self):
        return sum(k[0] for k in self._kv), self._kv[0][1]

    def set_batch_size_observer(self, fn) -> None:
        """`fn(batch_size)` per scheduler step of every replica (they arrive in bundles with the stats message)."""
        self._batch_observer = fn

    def stats(self) -> dict:
        return {"kv_used_blocks": sum(self._used), "replica_load": list(self.router.load)}

    def shutdown(self) -> None:
        self._stop = True
        for r in range(self.n):
            self._outq[r].put(("stop",))
            self._outq[r].put(None)
        for t in self._writers:
            t.join(timeout=5)
        for p in self._procs:
            p.join(timeout=5)
            if p.is_alive():
                p.terminate()
