"""Synthetic AgentVerse traffic: generator, replayer and writer in the reference's on-disk formats.

The reference ships no recorded traces; what it defines is the SHAPE of the traffic
(agents/agent_a/orchestrator.py:1859-2046: recruitment -> decision rounds -> N concurrent executions ->
evaluation -> final synthesis, N <= 5, closed loop) and the record format its tooling reads:
`response.json -> llm_requests[] {source, start_time_utc, llm_meta{prompt_tokens, completion_tokens}}`
(scripts/experiment/plot_results.py:674-686) and `logs/llm_calls.jsonl`
(agents/common/metrics_logger.py:27-72). This module generates seeded workflows of that shape
(BASELINE.json configs[2]), replays them against any engine with the
`generate(prompt_ids, params, request_id)` async-iterator interface, and writes the records back out
in those formats so the reference's plot / correlate scripts can consume a replay.
"""
from __future__ import annotations

import asyncio
import json
import random
import time
from dataclasses import dataclass, field
from datetime import datetime, timezone
from pathlib import Path
from typing import Dict, List, Optional, Sequence

from .scheduler import SamplingParams


@dataclass
class Call:
    source: str                 # "agent_a" (orchestrator) or "agent_b_<k>" (worker)
    stage: str
    prompt_ids: List[int]
    max_tokens: int


@dataclass
class Stage:
    name: str
    calls: List[Call]           # issued concurrently; the stage ends when all have answered


@dataclass
class Task:
    task_id: str
    stages: List[Stage]


@dataclass
class CallRecord:
    task_id: str
    request_id: str
    source: str
    stage: str
    start_time_utc: str
    start_s: float
    ttft_s: float
    latency_s: float
    prompt_tokens: int
    completion_tokens: int
    cached_prompt_tokens: int = 0


def _lognormal_int(rng: random.Random, median: float, sigma: float, lo: int, hi: int) -> int:
    import math

    return int(min(hi, max(lo, round(median * math.exp(rng.gauss(0.0, sigma))))))


def generate_tasks(num_tasks: int, seed: int = 3, vocab: int = 128000, max_prompt: int = 3500,
                   out_median: int = 220, out_cap: int = 512) -> List[Task]:
    """Seeded closed-loop workflows. Prompts of one task share the template + task text prefix; the
    executions additionally share the decision text (agents/agent_a/prompts.py:85-100)."""
    rng = random.Random(seed)

    def toks(n: int) -> List[int]:
        return [rng.randrange(vocab) for _ in range(n)]

    tasks: List[Task] = []
    for t in range(num_tasks):
        n_workers = rng.randint(1, 5)
        rounds = rng.randint(1, 3)
        base = toks(_lognormal_int(rng, 300, 0.3, 64, 800))               # template + task description

        def call(source: str, stage: str, extra: Sequence[int], cap: int = out_cap) -> Call:
            prompt = (base + list(extra))[:max_prompt]
            return Call(source, stage, prompt, min(cap, _lognormal_int(rng, out_median, 0.5, 16, cap)))

        stages = [Stage("recruitment", [call("agent_a", "recruitment", toks(_lognormal_int(rng, 100, 0.4, 16, 400)))])]
        decision: List[int] = []
        for r in range(rounds):                                           # vertical decision making
            proposal = toks(_lognormal_int(rng, 350, 0.5, 32, 1500))
            stages.append(Stage(f"decision_{r}_solver", [call("agent_b_0", "decision_solver", decision + proposal)]))
            if n_workers > 1:
                stages.append(Stage(f"decision_{r}_review", [
                    call(f"agent_b_{k}", "decision_review", decision + proposal + toks(40)) for k in range(1, n_workers)]))
            decision = (decision + proposal)[-1500:]
        stages.append(Stage("execution", [call(f"agent_b_{k}", "execution", decision + toks(60)) for k in range(n_workers)]))
        stages.append(Stage("evaluation", [call("agent_a", "evaluation", decision + toks(200 * n_workers))]))
        stages.append(Stage("final", [call("agent_a", "final", decision + toks(150 * n_workers), cap=out_cap)]))
        tasks.append(Task(f"task-{seed}-{t}", stages))
    return tasks


async def replay(engine, tasks: Sequence[Task], think_ms=(5.0, 20.0), seed: int = 3, temperature: float = 0.0,
                 ignore_eos: bool = True, id_prefix: str = "") -> List[CallRecord]:
    """Run every task concurrently (closed loop inside a task, think-time gap between stages)."""
    rng = random.Random(seed + 17)
    records: List[CallRecord] = []
    t_origin = time.monotonic()

    async def one_call(task: Task, idx: int, c: Call) -> None:
        rid = f"{id_prefix}{task.task_id}-{idx}"
        params = SamplingParams(temperature=temperature, max_tokens=c.max_tokens, ignore_eos=ignore_eos)
        t0 = time.monotonic()
        wall = datetime.now(timezone.utc).isoformat()
        ttft, n_out, cached = None, 0, 0
        async for out in engine.generate(c.prompt_ids, params, rid):
            if ttft is None:
                ttft = time.monotonic() - t0
            n_out, cached = len(out.token_ids), getattr(out, "num_cached_tokens", 0)
        records.append(CallRecord(task.task_id, rid, c.source, c.stage, wall, t0 - t_origin, ttft or 0.0,
                                  time.monotonic() - t0, len(c.prompt_ids), n_out, cached))

    async def one_task(task: Task) -> None:
        k = 0
        for st in task.stages:
            await asyncio.gather(*[one_call(task, k + i, c) for i, c in enumerate(st.calls)])
            k += len(st.calls)
            await asyncio.sleep(rng.uniform(*think_ms) / 1e3)

    await asyncio.gather(*[one_task(t) for t in tasks])
    return sorted(records, key=lambda r: r.start_s)


def write_reference_formats(records: Sequence[CallRecord], out_dir: Path) -> None:
    """`<out_dir>/<task_id>/response.json` (one per task) + `<out_dir>/llm_calls.jsonl`."""
    out_dir = Path(out_dir)
    out_dir.mkdir(parents=True, exist_ok=True)
    by_task: Dict[str, List[CallRecord]] = {}
    for r in records:
        by_task.setdefault(r.task_id, []).append(r)
    for task_id, recs in by_task.items():
        d = out_dir / task_id
        d.mkdir(exist_ok=True)
        body = {"task_id": task_id, "llm_requests": [
            {"source": r.source, "stage": r.stage, "request_id": r.request_id, "start_time_utc": r.start_time_utc,
             "llm_meta": {"request_id": r.request_id, "prompt_tokens": r.prompt_tokens,
                          "completion_tokens": r.completion_tokens,
                          "total_tokens": r.prompt_tokens + r.completion_tokens,
                          "latency_ms": int(r.latency_s * 1000), "queue_wait_s": round(r.ttft_s, 4)}}
            for r in recs]}
        (d / "response.json").write_text(json.dumps(body, indent=1))
    with open(out_dir / "llm_calls.jsonl", "w") as f:
        for r in records:
            f.write(json.dumps({"timestamp": r.start_time_utc, "task_id": r.task_id, "agent": r.source,
                                "request_id": r.request_id, "prompt_tokens": r.prompt_tokens,
                                "completion_tokens": r.completion_tokens,
                                "total_tokens": r.prompt_tokens + r.completion_tokens,
                                "latency_ms": int(r.latency_s * 1000)}) + "\n")


def load_reference_run(run_dir: Path) -> List[dict]:
    """Read `llm_requests` from every response.json under a reference `data/runs/<run>/` tree, sorted by
    start time: the arrival process a user's real run produced (for replay with synthetic ids)."""
    reqs: List[dict] = []
    for p in sorted(Path(run_dir).rglob("response.json")):
        body = json.loads(p.read_text())
        for r in body.get("llm_requests", []):
            meta = r.get("llm_meta") or {}
            if r.get("start_time_utc") and meta.get("prompt_tokens") is not None:
                reqs.append({"source": r.get("source", "agent_a"), "start_time_utc": r["start_time_utc"],
                             "prompt_tokens": int(meta["prompt_tokens"]),
                             "completion_tokens": int(meta.get("completion_tokens") or 0)})
    return sorted(reqs, key=lambda r: r["start_time_utc"])


def summarize(records: Sequence[CallRecord], wall_s: float) -> dict:
    import statistics

    out = sum(r.completion_tokens for r in records)
    ttft = sorted(r.ttft_s for r in records)
    return {"requests": len(records), "output_tokens": out, "prompt_tokens": sum(r.prompt_tokens for r in records),
            "cached_prompt_tokens": sum(r.cached_prompt_tokens for r in records),
            "tokens_per_s": out / wall_s if wall_s > 0 else 0.0,
            "ttft_p50_ms": 1e3 * statistics.median(ttft) if ttft else None,
            "ttft_p95_ms": 1e3 * ttft[int(0.95 * (len(ttft) - 1))] if ttft else None, "wall_s": wall_s}
