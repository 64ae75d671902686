"""HTTP / metrics / tracing contract of the drop-in server, checked on CPU with a fake engine, and
differentially against golden facts extracted from the reference (tests/golden/serve_contract.json,
minted by oracle/gen_contract_golden.py which imports the reference's llm/serve_llm.py)."""
import asyncio
import json
from pathlib import Path

import pytest
from aiohttp.test_utils import TestClient, TestServer

GOLD = json.loads((Path(__file__).parent / "golden" / "serve_contract.json").read_text())


@pytest.fixture()
def served(monkeypatch):
    from llm import serve_llm as S
    from llm.backend import B200Backend
    from tests.fakes import FakeAsyncEngine

    eng = FakeAsyncEngine()
    backend = B200Backend("random:tiny", None, None, None, None, None, tracer=S._tracer, engine=eng)
    monkeypatch.setattr(S, "_backend", backend)
    loop = asyncio.new_event_loop()
    client = TestClient(TestServer(S.create_app()), loop=loop)
    loop.run_until_complete(client.start_server())
    yield S, eng, client, loop
    loop.run_until_complete(client.close())
    loop.close()


def _post(client, loop, path, **kw):
    async def go():
        r = await client.post(path, **kw)
        return r.status, await r.json(), r.headers
    return loop.run_until_complete(go())


def test_routes_match_reference(served):
    S, _, _, _ = served
    ours = sorted((r.method, r.resource.canonical) for r in S.create_app().router.routes() if r.method != "HEAD")
    for method, path in GOLD["routes"]:
        assert (method, path) in ours, (method, path)


@pytest.mark.parametrize("path", ["/chat", "/completion", "/generate"])
def test_chat_response_shape(served, path):
    S, eng, client, loop = served
    status, body, _ = _post(client, loop, path, json={"prompt": "hello", "max_tokens": 3},
                            headers={"X-Request-ID": "abc12345",
                                     "traceparent": "00-0af7651916cd43dd8448eb211c80319c-b7ad6b7169203331-01"})
    assert status == 200
    assert set(body) == {"output", "meta"} and isinstance(body["output"], str) and len(body["output"]) == 3
    meta = body["meta"]
    assert set(meta) == set(GOLD["meta_keys"])
    assert meta["request_id"] == "abc12345"
    assert isinstance(meta["latency_ms"], int) and isinstance(meta["queue_wait_s"], float)
    assert meta["completion_tokens"] == 3 and meta["total_tokens"] == meta["prompt_tokens"] + 3
    otel = meta["otel"]
    assert otel["trace_id"] == "0af7651916cd43dd8448eb211c80319c" and len(otel["span_id"]) == 16
    assert otel["attributes"]["app.path"] == path and otel["attributes"]["app.request_id"] == "abc12345"
    assert otel["attributes"]["app.chat_template_applied"] is True
    assert eng.seen[-1]["params"].max_tokens == 3 and eng.seen[-1]["params"].temperature == 0.2


def test_chat_template_is_the_reference_fallback_string(served):
    S, eng, client, loop = served
    _post(client, loop, "/chat", json={"prompt": GOLD["template_case"]["prompt"],
                                       "system_prompt": GOLD["template_case"]["system_prompt"]})
    tok = S._backend._tokenizer
    assert tok.decode(eng.seen[-1]["ids"]) == GOLD["template_case"]["formatted"]
    _post(client, loop, "/chat", json={"prompt": GOLD["template_case"]["prompt"]})
    assert tok.decode(eng.seen[-1]["ids"]) == GOLD["template_case"]["formatted_default_system"]
    # a raw prompt reaches the model as vLLM would feed it: <|begin_of_text|> first (exactly once), then the text
    _post(client, loop, "/chat", json={"input": "raw", "skip_chat_template": True})
    ids = eng.seen[-1]["ids"]
    assert ids[0] == tok.bos_token_id and tok.decode(ids[1:]) == "raw"
    _post(client, loop, "/chat", json={"input": "<|begin_of_text|>raw", "skip_chat_template": True})
    assert eng.seen[-1]["ids"] == ids


def test_error_bodies_and_status_codes(served):
    S, eng, client, loop = served
    st, body, _ = _post(client, loop, "/chat", data=b"{not json", headers={"Content-Type": "application/json"})
    assert (st, body) == (400, {"error": "Invalid JSON"})
    for bad in ({}, {"prompt": ""}, {"prompt": 7}, {"input": None}):
        st, body, _ = _post(client, loop, "/chat", json=bad)
        assert (st, body) == (400, {"error": "Missing 'prompt' field"})
    eng.fail = True
    st, body, _ = _post(client, loop, "/chat", json={"prompt": "x"})
    assert st == 500 and body == {"error": "Generation failed: boom"}
    eng.fail = False
    assert S._Traffic.inflight == 0  # every exit path released its slot
    import llm.serve_llm as mod

    keep, mod._backend = mod._backend, None
    try:
        st, body, _ = _post(client, loop, "/chat", json={"prompt": "x"})
        assert (st, body) == (503, {"error": "Backend not initialized"})
    finally:
        mod._backend = keep


def test_max_tokens_coercion(served):
    S, eng, client, loop = served
    _post(client, loop, "/chat", json={"prompt": "x", "max_tokens": "4"})
    assert eng.seen[-1]["params"].max_tokens == 4
    _post(client, loop, "/chat", json={"prompt": "x", "max_tokens": "many"})
    assert eng.seen[-1]["params"].max_tokens == S.LLM_MAX_TOKENS
    _post(client, loop, "/chat", json={"prompt": "x", "max_new_tokens": 2})  # hf_cpu_server.py:81 alias
    assert eng.seen[-1]["params"].max_tokens == 2


def test_health_and_metrics_exposition(served):
    S, eng, client, loop = served
    S.export_static_gauges(S._backend, 0.9)
    _post(client, loop, "/chat", json={"prompt": "hello"})

    async def get(path):
        r = await client.get(path)
        return r.status, await r.text(), r.headers
    for p in ("/health", "/ready", "/live"):
        st, text, _ = loop.run_until_complete(get(p))
        assert st == 200 and json.loads(text) == {"status": "ok"}
    st, text, hdr = loop.run_until_complete(get("/metrics"))
    assert st == 200 and hdr["Content-Type"] == GOLD["metrics_content_type"]
    from prometheus_client.parser import text_string_to_metric_families

    fams = {f.name: f for f in text_string_to_metric_families(text)}
    for name, spec in GOLD["metric_families"].items():
        assert name in fams, name
        assert fams[name].type == spec["type"], name
        assert fams[name].documentation == spec["help"], name
        if spec["type"] == "histogram":
            les = sorted({float(s.labels["le"]) for s in fams[name].samples if s.name.endswith("_bucket")})
            assert les == sorted(spec["buckets"] + [float("inf")]), name
    val = {s.name: s.value for f in fams.values() for s in f.samples if not s.labels}
    assert val["llm_kv_cache_num_gpu_blocks"] == 1000 and val["llm_kv_cache_total_tokens"] == 16000
    assert val["llm_kv_cache_est_max_concurrency_at_max_model_len"] == 16000 // 4096
    assert abs(val["llm_computed_max_concurrency"] - 16000 / 4096) < 1e-9
    assert val["llm_config_max_num_seqs"] == 12 and val["llm_config_gpu_memory_utilization"] == 0.9
    assert val["llm_inflight_requests"] == 0
    assert val["llm_prompt_tokens_total"] > 0 and val["llm_completion_tokens_total"] > 0
    # every llm_* series the Grafana dashboard queries exists in our exposition
    series = {s.name for f in fams.values() for s in f.samples}
    for name in GOLD["grafana_series"]:
        assert name in series or name in fams, name


def test_spans_have_reference_names_and_parentage(served):
    from opentelemetry import trace
    from opentelemetry.sdk.trace.export import SimpleSpanProcessor
    from opentelemetry.sdk.trace.export.in_memory_span_exporter import InMemorySpanExporter

    S, eng, client, loop = served
    exp = InMemorySpanExporter()
    trace.get_tracer_provider().add_span_processor(SimpleSpanProcessor(exp))
    _post(client, loop, "/chat", json={"prompt": "hello"})
    spans = {s.name: s for s in exp.get_finished_spans()}
    assert set(GOLD["span_names"]) <= set(spans)
    root = spans["llm.handle_request"]
    assert root.kind.name == "SERVER"
    for child in ("llm.time_to_first_token", "llm.generate"):
        assert spans[child].parent.span_id == root.context.span_id
    assert "llm_ttft_seconds" in spans["llm.time_to_first_token"].attributes
    assert "llm.generate_ms" in spans["llm.generate"].attributes
    for key in ("app.path", "app.request_id", "app.prompt_length", "app.formatted_prompt_length",
                "app.chat_template_applied", "app.prompt_truncated", "llm.prompt_tokens", "llm.completion_tokens",
                "llm.total_tokens"):
        assert key in root.attributes, key


def test_prompt_truncation_keeps_head(served, monkeypatch):
    S, eng, client, loop = served
    monkeypatch.setattr(S, "LLM_MAX_MODEL_LEN", 300)
    monkeypatch.setattr(S, "LLM_PROMPT_SAFETY_MARGIN_TOKENS", 28)
    st, body, _ = _post(client, loop, "/chat", json={"prompt": "z" * 1000, "max_tokens": 72, "skip_chat_template": True})
    assert st == 200 and len(eng.seen[-1]["ids"]) == 201  # 300 - 72 - 28 text tokens after <|begin_of_text|>
    assert eng.seen[-1]["ids"][1:] == [ord("z")] * 200
    assert body["meta"]["otel"]["attributes"]["app.prompt_truncated"] is True
    assert body["meta"]["otel"]["attributes"]["app.prompt_truncated_tokens"] == 800
    assert body["meta"]["prompt_tokens"] == 200


def test_openai_chat_completions_shape(served):
    S, eng, client, loop = served
    msgs = [{"role": "system", "content": "be brief"}, {"role": "user", "content": [{"type": "text", "text": "hi"}]}]
    st, body, _ = _post(client, loop, "/v1/chat/completions", json={"model": "m", "messages": msgs, "max_tokens": 2})
    assert st == 200 and body["object"] == "chat.completion" and body["model"] == "m"
    assert body["id"].startswith("chatcmpl-") and len(body["id"]) == len("chatcmpl-") + 24
    ch = body["choices"][0]
    assert ch["index"] == 0 and ch["message"]["role"] == "assistant" and len(ch["message"]["content"]) == 2
    assert ch["finish_reason"] == "length"
    assert body["usage"]["completion_tokens"] == 2 and body["usage"]["total_tokens"] == body["usage"]["prompt_tokens"] + 2
    # default prompt == what the reference proxy + /chat would have built
    tok = S._backend._tokenizer
    assert GOLD["flatten_case"]["flattened"] in tok.decode(eng.seen[-1]["ids"])
    st, body, _ = _post(client, loop, "/v1/chat/completions", json={"messages": []})
    assert st == 400 and body == {"error": "Field 'messages' must be a non-empty list"}
    st, body, _ = _post(client, loop, "/v1/completions", json={"prompt": "abc", "max_tokens": 1})
    assert st == 200 and body["object"] == "text_completion" and tok.decode(eng.seen[-1]["ids"][1:]) == "abc"


def test_chat_streaming_sse(served):
    S, eng, client, loop = served

    async def go():
        r = await client.post("/chat", json={"prompt": "hello", "max_tokens": 4, "stream": True})
        assert r.status == 200 and r.headers["Content-Type"].startswith("text/event-stream")
        return (await r.read()).decode()
    body = loop.run_until_complete(go())
    events = [e[len("data: "):] for e in body.strip().split("\n\n")]
    assert events[-1] == "[DONE]"
    deltas = [json.loads(e)["delta"] for e in events[:-2]]
    final = json.loads(events[-2])
    assert len(deltas) == 4 and "".join(deltas) == final["output"]
    assert final["meta"]["completion_tokens"] == 4 and S._Traffic.inflight == 0


def test_cli_and_env_precedence(monkeypatch):
    from llm import serve_llm as S

    captured = {}

    async def fake_run(*a):
        captured["args"] = a
    monkeypatch.setattr(S, "run_async_server", fake_run)
    monkeypatch.setattr(S, "LLM_MAX_NUM_SEQS", "7")
    monkeypatch.setattr(S, "LLM_DTYPE", "float16")
    monkeypatch.setattr(S, "LLM_GPU_MEMORY_UTILIZATION", "0.5")
    S.main(["--model", "random:tiny", "--port", "8123", "--max-num-seqs", "3", "--max-model-len", "512"])
    host, port, model, mml, dtype, seqs, batched, util = captured["args"]
    assert (host, port, model, mml, dtype, seqs, batched, util) == ("0.0.0.0", 8123, "random:tiny", 512, "float16", 3,
                                                                   None, 0.5)
    with pytest.raises(SystemExit) as e:
        S.main(["--tp-size", "8"])  # unknown flags stay an argparse error (exit 2), as in the reference
    assert e.value.code == 2


def test_cli_parallelism_flags_are_additive_and_set_the_env(monkeypatch):
    """--tensor-parallel-size / --data-parallel-size do not exist in the reference CLI (llm/serve_llm.py:1049-1104
    is single-GPU); here they are additive switches that only select how B200Backend builds its engine."""
    import os

    from llm import serve_llm as S

    async def fake_run(*a):
        pass
    monkeypatch.setattr(S, "run_async_server", fake_run)
    monkeypatch.delenv("LLM_TP_SIZE", raising=False)
    monkeypatch.delenv("LLM_DP_SIZE", raising=False)
    S.main(["--model", "random:tiny"])
    assert "LLM_TP_SIZE" not in os.environ and "LLM_DP_SIZE" not in os.environ
    S.main(["--model", "random:70b:0", "--tensor-parallel-size", "8"])
    assert os.environ["LLM_TP_SIZE"] == "8" and "LLM_DP_SIZE" not in os.environ
    monkeypatch.delenv("LLM_TP_SIZE", raising=False)
    S.main(["--model", "random:8b:0", "--data-parallel-size", "4"])
    assert os.environ["LLM_DP_SIZE"] == "4"
    monkeypatch.delenv("LLM_DP_SIZE", raising=False)


def test_backend_rejects_tp_and_dp_together(monkeypatch):
    from llm.backend import B200Backend

    monkeypatch.setenv("LLM_TP_SIZE", "2")
    monkeypatch.setenv("LLM_DP_SIZE", "2")
    with pytest.raises(ValueError, match="cannot both"):
        B200Backend("random:tiny", 256, None, 4, 64, 0.5)
