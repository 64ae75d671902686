"""Tensor-parallel worker group for the server (the 70B model shards over one box).

`bench.py --workload tp70b` runs the TP engine under torchrun, every rank feeding itself the same requests.
A server has ONE front door, so here rank 0 is the engine thread of the HTTP process and ranks 1..N-1 are
follower processes, one per GPU. The ranks must make identical scheduling decisions (the kernels exchange
partial sums row by row and every rank samples the same token), and the scheduler is deterministic given
its inputs, so the only thing that has to cross processes on the host side is the request stream:

    rank 0, once per engine-loop iteration:  send(list of add/abort messages)  ->  every follower
    follower:                                 apply them in order, then step iff it has work

exactly the sequence of calls rank 0 makes itself. No tokens travel back: every rank computes them. Device
work stays in lockstep through the collectives inside the step; the pipes only need to preserve order.
"""
from __future__ import annotations

import dataclasses
import multiprocessing as mp
import os
import socket
from typing import List

from .config import EngineConfig


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _init_rank(rank: int, world: int, port: int):
    import torch
    import torch.distributed as dist

    os.environ.setdefault("NCCL_DEBUG", "WARN")
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                            device_id=torch.device("cuda", rank))
    return dist


def apply_messages(eng, msgs) -> List[tuple]:
    """Apply add/abort messages to an engine; returns [(request_id, exception)] for rejected requests.
    Used by rank 0 and by the followers, so both sides see the same accept/reject decisions."""
    errors = []
    for msg in msgs:
        if msg[0] == "add":
            try:
                eng.add_request(msg[1], msg[2], msg[3])
            except Exception as exc:
                errors.append((msg[1], exc))
        elif msg[0] == "abort":
            eng.abort(msg[1])
    return errors


def _follower(rank: int, world: int, port: int, model: str, ecfg: EngineConfig, conn) -> None:
    try:
        dist = _init_rank(rank, world, port)
        from .engine import build_engine

        eng = build_engine(model, dataclasses.replace(ecfg, device=rank, tp_size=world, tp_rank=rank))
        conn.send(("ready", rank, eng.runner.kv_cache_info()))
    except BaseException as exc:
        conn.send(("fatal", rank, repr(exc)))
        return
    try:
        while True:
            msg = conn.recv()
            if msg[0] == "stop":
                break
            apply_messages(eng, msg[1])
            if eng.has_work():
                eng.step()
    except BaseException as exc:   # device fault, lost pipe: tell rank 0 (it checks before every broadcast)
        try:
            conn.send(("fatal", rank, repr(exc)))
        except (OSError, BrokenPipeError):
            pass
        raise
    finally:
        eng.runner.close()
        dist.destroy_process_group()


class TPGroup:
    """Spawns the followers, joins the process group as rank 0 and builds rank 0's engine (in the caller's
    process). `broadcast(msgs)` must be called once per engine-loop iteration, before rank 0 applies `msgs`."""

    def __init__(self, model: str, ecfg: EngineConfig, tp: int):
        from .engine import build_engine

        if tp < 2:
            raise ValueError("TPGroup needs tp >= 2")
        self.tp = tp
        port = _free_port()
        ctx = mp.get_context("spawn")
        self._conns, self._procs = [], []
        for r in range(1, tp):
            parent, child = ctx.Pipe(duplex=True)
            p = ctx.Process(target=_follower, args=(r, tp, port, model, ecfg, child), daemon=True, name=f"atx-tp-{r}")
            p.start()
            self._conns.append(parent)
            self._procs.append(p)
        try:
            self._dist = _init_rank(0, tp, port)
            self.engine = build_engine(model, dataclasses.replace(ecfg, device=0, tp_size=tp, tp_rank=0))
            for r, c in enumerate(self._conns, start=1):
                msg = c.recv()
                if msg[0] != "ready":
                    raise RuntimeError(f"tensor-parallel rank {r} failed to start: {msg[2]}")
                if msg[2] != self.engine.runner.kv_cache_info():
                    raise RuntimeError(f"rank {r} sized its KV pool differently: {msg[2]}")
        except BaseException:
            for p in self._procs:
                if p.is_alive():
                    p.kill()
            raise

    def broadcast(self, msgs) -> None:
        """Raises RuntimeError when a follower has died or reported a failure: the engine thread then fails
        every waiting request (AsyncEngine._run) instead of launching a step whose peers will never answer."""
        item = ("step", list(msgs))
        for r, (c, p) in enumerate(zip(self._conns, self._procs), start=1):
            try:
                if c.poll(0):
                    msg = c.recv()
                    if msg and msg[0] == "fatal":
                        raise RuntimeError(f"tensor-parallel rank {r} failed: {msg[2]}")
            except (EOFError, OSError) as exc:
                raise RuntimeError(f"tensor-parallel rank {r} closed its pipe") from exc
            if not p.is_alive():
                raise RuntimeError(f"tensor-parallel rank {r} exited (code {p.exitcode})")
            try:
                c.send(item)
            except (OSError, BrokenPipeError) as exc:
                raise RuntimeError(f"tensor-parallel rank {r} is gone: {exc!r}") from exc

    def shutdown(self) -> None:
        for c in self._conns:
            try:
                c.send(("stop",))
            except (OSError, BrokenPipeError):
                pass
        self.engine.runner.close()
        try:
            self._dist.destroy_process_group()   # the followers are doing the same right now
        except Exception:
            pass
        for p in self._procs:
            p.join(timeout=20)
            if p.is_alive():
                p.kill()
