"""Synthetic AgentVerse trace generator / replayer / reference-format writer (CPU, fake engine)."""
import asyncio
import json

from atx import trace as TR
from tests.fakes import FakeAsyncEngine


def test_generator_is_seeded_and_has_the_workflow_shape():
    a, b = TR.generate_tasks(5, seed=3), TR.generate_tasks(5, seed=3)
    assert [[len(c.prompt_ids) for s in t.stages for c in s.calls] for t in a] == \
           [[len(c.prompt_ids) for s in t.stages for c in s.calls] for t in b]
    assert TR.generate_tasks(5, seed=4)[0].stages[0].calls[0].prompt_ids != a[0].stages[0].calls[0].prompt_ids
    for t in a:
        names = [s.name for s in t.stages]
        assert names[0] == "recruitment" and names[-3:] == ["execution", "evaluation", "final"]
        ex = t.stages[-3].calls
        assert 1 <= len(ex) <= 5 and all(c.source.startswith("agent_b_") for c in ex)
        assert all(len(s.calls) <= 5 for s in t.stages)
        for s in t.stages:
            for c in s.calls:
                assert 64 <= len(c.prompt_ids) <= 3500 and 16 <= c.max_tokens <= 512
        # concurrent executions share template + task + decision text (prefix-cache friendly)
        if len(ex) > 1:
            n = min(len(c.prompt_ids) for c in ex) - 60
            assert ex[0].prompt_ids[:n] == ex[1].prompt_ids[:n] and n > 64


def test_replay_and_reference_formats(tmp_path):
    tasks = TR.generate_tasks(3, seed=9)
    eng = FakeAsyncEngine()
    recs = asyncio.run(TR.replay(eng, tasks, think_ms=(0.1, 0.2)))
    n_calls = sum(len(s.calls) for t in tasks for s in t.stages)
    assert len(recs) == n_calls == len(eng.seen)
    assert all(r.completion_tokens == min(5, c_max) for r, c_max in zip(recs, [5] * len(recs)))
    # closed loop: a task's stage never starts before its previous stage ended
    for t in tasks:
        mine = [r for r in recs if r.task_id == t.task_id]
        k = 0
        prev_end = 0.0
        for st in t.stages:
            grp = [r for r in mine if r.stage == st.calls[0].stage and r.request_id.split("-")[-1] in
                   {str(i) for i in range(k, k + len(st.calls))}]
            assert len(grp) == len(st.calls)
            assert min(r.start_s for r in grp) >= prev_end - 1e-6
            prev_end = max(r.start_s + r.latency_s for r in grp)
            k += len(st.calls)
    TR.write_reference_formats(recs, tmp_path)
    body = json.loads((tmp_path / tasks[0].task_id / "response.json").read_text())
    req = body["llm_requests"][0]
    assert {"source", "start_time_utc", "llm_meta"} <= set(req)
    assert {"prompt_tokens", "completion_tokens", "total_tokens", "latency_ms", "request_id"} <= set(req["llm_meta"])
    lines = (tmp_path / "llm_calls.jsonl").read_text().splitlines()
    assert len(lines) == n_calls and json.loads(lines[0])["prompt_tokens"] > 0
    back = TR.load_reference_run(tmp_path)
    assert len(back) == n_calls and back == sorted(back, key=lambda r: r["start_time_utc"])
    s = TR.summarize(recs, 2.0)
    assert s["requests"] == n_calls and s["tokens_per_s"] == s["output_tokens"] / 2.0 and s["ttft_p50_ms"] >= 0
