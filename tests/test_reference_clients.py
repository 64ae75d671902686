"""The reference's OWN clients, imported unmodified from /root/reference, against the drop-in server
(SURVEY.md section 4 i): agent B's and agent A's `call_llm`, the orchestrator's `_call_llm`, the health
check, the MCP-Universe OpenAI proxy pointed at /chat, the per-call metrics logger reading `meta` — and the
reference's literal `HFRequestHandler` (llm/hf_cpu_server.py:54-94, stub pipeline) next to ours, request for
request. CPU only (fake engine). Skipped where /root/reference does not exist (the GPU box).
"""
import asyncio
import importlib
import importlib.util
import json
import os
import sys
import threading
from pathlib import Path

import pytest
from aiohttp import web

REF = Path("/root/reference")
pytestmark = pytest.mark.skipif(not (REF / "agents" / "agent_b" / "main.py").exists(),
                                reason="/root/reference is not present on this machine")


class _Served:
    """An aiohttp app on a real 127.0.0.1 port, its loop running in a background thread (the reference clients
    are synchronous httpx callers)."""

    def __init__(self, app: web.Application):
        self.loop = asyncio.new_event_loop()
        self.port = None
        ready = threading.Event()

        async def start():
            self.runner = web.AppRunner(app)
            await self.runner.setup()
            site = web.TCPSite(self.runner, "127.0.0.1", 0)
            await site.start()
            self.port = site._server.sockets[0].getsockname()[1]
            ready.set()

        def run():
            asyncio.set_event_loop(self.loop)
            self.loop.run_until_complete(start())
            self.loop.run_forever()

        self.thread = threading.Thread(target=run, daemon=True)
        self.thread.start()
        assert ready.wait(20)

    def url(self, path: str) -> str:
        return f"http://127.0.0.1:{self.port}{path}"

    def close(self):
        async def stop():
            await self.runner.cleanup()
        asyncio.run_coroutine_threadsafe(stop(), self.loop).result(20)
        self.loop.call_soon_threadsafe(self.loop.stop)
        self.thread.join(20)


@pytest.fixture()
def llm_server(monkeypatch):
    from llm import serve_llm as S
    from llm.backend import B200Backend
    from tests.fakes import FakeAsyncEngine

    eng = FakeAsyncEngine()
    monkeypatch.setattr(S, "_backend", B200Backend("random:tiny", None, None, None, None, None, tracer=S._tracer,
                                                   engine=eng))
    srv = _Served(S.create_app())
    yield srv, eng
    srv.close()


def _import_reference(name: str, env: dict):
    """Import a reference module fresh, with the env it reads at import time (LLM_SERVER_URL ...). The
    reference's top-level directories (`agents`, `tools`) are namespace packages without __init__.py; an
    unrelated regular package of the same name in site-packages would win the import, so the top-level name
    is bound to the reference directory for the duration of the import."""
    import types

    top = name.split(".")[0]
    old_env = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    saved = {m: sys.modules.pop(m) for m in list(sys.modules) if m == top or m.startswith(top + ".")}
    sys.path.insert(0, str(REF))
    try:
        if (REF / top).is_dir():
            pkg = types.ModuleType(top)
            pkg.__path__ = [str(REF / top)]
            sys.modules[top] = pkg
        mod = importlib.import_module(name)
        mod._ref_modules = {m: sys.modules[m] for m in sys.modules if m == top or m.startswith(top + ".")}
        return mod
    finally:
        sys.path.remove(str(REF))
        for m in [m for m in sys.modules if m == top or m.startswith(top + ".")]:
            del sys.modules[m]
        sys.modules.update(saved)
        for k, v in old_env.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


META_KEYS = {"request_id", "latency_ms", "queue_wait_s", "prompt_tokens", "completion_tokens", "total_tokens", "otel"}


@pytest.mark.parametrize("module", ["agents.agent_b.main", "agents.agent_a.main"])
def test_agents_call_llm_unmodified(llm_server, module, tmp_path, monkeypatch):
    srv, eng = llm_server
    monkeypatch.chdir(tmp_path)          # the agents' telemetry logger writes under ./logs
    mod = _import_reference(module, {"LLM_SERVER_URL": srv.url("/chat")})
    hdrs = {"X-Request-ID": "feedc0de", "X-Task-ID": "11111111-2222-3333-4444-555555555555",
            "traceparent": "00-0af7651916cd43dd8448eb211c80319c-b7ad6b7169203331-01"}
    out, meta = mod.call_llm("Summarise the plan in one line.", headers=hdrs)
    assert isinstance(out, str) and out
    assert set(meta) == META_KEYS and meta["request_id"] == "feedc0de"
    assert meta["otel"]["trace_id"] == "0af7651916cd43dd8448eb211c80319c"     # the caller's trace continues
    assert meta["total_tokens"] == meta["prompt_tokens"] + meta["completion_tokens"]
    assert eng.seen[-1]["params"].temperature == 0.2 and eng.seen[-1]["params"].max_tokens == 512
    out2, meta2 = mod.call_llm("no headers at all")
    assert out2 and len(meta2["request_id"]) == 8


def test_orchestrator_call_llm_unmodified(llm_server, tmp_path, monkeypatch):
    """agents/agent_a/orchestrator.py:217-258: the only caller that sends max_tokens and reads meta.otel."""
    srv, eng = llm_server
    monkeypatch.chdir(tmp_path)
    orch = _import_reference("agents.agent_a.orchestrator", {"LLM_SERVER_URL": srv.url("/chat"),
                                                             "OTEL_EXPORTER_OTLP_ENDPOINT": "http://127.0.0.1:9/x"})
    logger = orch._ref_modules["agents.common.telemetry"].TelemetryLogger(agent_id="AgentA", scenario="test")
    o = orch.AgentVerseOrchestrator(logger)
    out, meta = o._call_llm("Recruit experts for: build a CLI.", headers={"X-Request-ID": "0badf00d"}, max_tokens=3)
    assert isinstance(out, str) and len(out) == 3
    assert eng.seen[-1]["params"].max_tokens == 3
    assert set(meta["llm_backend"]) == META_KEYS
    assert meta["otel"]["llm_backend"] == meta["llm_backend"]["otel"] and "trace_id" in meta["otel"]["llm_backend"]


def test_health_check_unmodified(llm_server):
    srv, _ = llm_server
    sys.path.insert(0, str(REF / "scripts" / "monitoring"))
    try:
        sys.modules.pop("health_check", None)
        hc = importlib.import_module("health_check")
    finally:
        sys.path.remove(str(REF / "scripts" / "monitoring"))
    assert hc.check_llm_server(srv.url("/chat")) == (True, None)
    assert hc.check_http_endpoint(srv.url("/health")) == (True, None)          # deploy.sh wait_for_llm polls this
    ok, err = hc.check_metrics_endpoint(srv.url("/metrics"), ["llm_requests_total", "llm_inflight_requests",
                                                               "llm_queue_wait_seconds_bucket"])
    assert ok, err


def test_openai_proxy_unmodified_forwards_to_chat(llm_server):
    """tools/mcp_universe/openai_proxy.py (the :8110 shim) keeps working against /chat."""
    import httpx

    srv, eng = llm_server
    proxy = _import_reference("tools.mcp_universe.openai_proxy", {})
    shim = _Served(proxy.create_app(srv.url("/chat")))
    try:
        msgs = [{"role": "system", "content": "be brief"}, {"role": "user", "content": [{"type": "text", "text": "hi"}]}]
        r = httpx.post(shim.url("/v1/chat/completions"), json={"model": "m", "messages": msgs, "max_tokens": 4}, timeout=30)
        assert r.status_code == 200
        body = r.json()
        assert body["object"] == "chat.completion" and body["model"] == "m"
        content = body["choices"][0]["message"]["content"]
        assert len(content) == 4 and not content.startswith("[Proxy]")
        assert eng.seen[-1]["params"].max_tokens == 4
        # what reached the engine is the proxy's flattened prompt inside the Llama-3 template
        from llm import serve_llm as S

        text = S._backend._tokenizer.decode(eng.seen[-1]["ids"])
        assert proxy._flatten_messages(msgs) in text
        # ... and the native route on :8000 answers the same request with the same content
        r2 = httpx.post(srv.url("/v1/chat/completions"), json={"model": "m", "messages": msgs, "max_tokens": 4}, timeout=30)
        assert r2.status_code == 200 and r2.json()["choices"][0]["message"]["content"] == content
    finally:
        shim.close()


def test_metrics_logger_finds_every_meta_key_it_reads(llm_server, tmp_path):
    """agents/common/metrics_logger.py:48-65 reads request_id / prompt_tokens / completion_tokens / total_tokens /
    latency_ms out of `meta`: all present and non-null from this backend."""
    import httpx

    srv, _ = llm_server
    ml = _import_reference("agents.common.metrics_logger", {})
    meta = httpx.post(srv.url("/chat"), json={"prompt": "hello"}, timeout=30).json()["meta"]
    lg = ml.MetricsLogger(log_dir=str(tmp_path))
    cid = lg.log_call(task_id="t", agent_id="AgentB", call_type="root", timestamp_start="a", timestamp_end="b",
                      http_status=200, llm_meta=meta)
    rec = json.loads((tmp_path / "llm_calls.jsonl").read_text().splitlines()[-1])
    assert cid == meta["request_id"] == rec["call_id"]
    for k in ("prompt_tokens", "completion_tokens", "total_tokens", "latency_ms"):
        assert isinstance(rec[k], int), k


# ------------------------------------------------------------------------- llm/hf_cpu_server.py, side by side
class _StubPipeline:
    """What transformers' text-generation pipeline returns to hf_cpu_server.py:91-92: full text."""

    def __init__(self):
        self.calls = []

    def __call__(self, prompt, **kw):
        self.calls.append((prompt, kw))
        return [{"generated_text": prompt + "abcde"[:kw["max_new_tokens"]]}]


@pytest.fixture()
def hf_pair(monkeypatch):
    """(reference handler on a real port with a stub pipeline, our handler on a real port with the fake engine)."""
    from http.server import HTTPServer

    spec = importlib.util.spec_from_file_location("ref_hf_cpu_server", REF / "llm" / "hf_cpu_server.py")
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    stub = _StubPipeline()
    ref._PIPELINE = stub
    httpd = HTTPServer(("127.0.0.1", 0), ref.HFRequestHandler)
    t = threading.Thread(target=httpd.serve_forever, daemon=True)
    t.start()

    from llm import hf_cpu_server as ours
    from llm.backend import B200Backend
    from tests.fakes import FakeAsyncEngine

    eng = FakeAsyncEngine()
    monkeypatch.setattr(ours, "_backend", B200Backend("random:tiny", None, None, None, None, None, engine=eng))
    srv = _Served(ours.create_app())
    yield f"http://127.0.0.1:{httpd.server_address[1]}", srv.url(""), stub, eng, ref
    srv.close()
    httpd.shutdown()
    httpd.server_close()


def test_hf_cpu_server_contract_side_by_side_with_the_reference_handler(hf_pair):
    import httpx

    ref_url, our_url, stub, eng, ref = hf_pair

    def both(method, path, **kw):
        a = httpx.request(method, ref_url + path, timeout=30, **kw)
        b = httpx.request(method, our_url + path, timeout=30, **kw)
        return a, b

    # error paths: same status, same body
    for path, kw in [("/chat", dict(content=b"{not json")), ("/chat", dict(json={})), ("/chat", dict(json={"prompt": ""})),
                     ("/chat", dict(json={"prompt": 5})), ("/nope", dict(json={"prompt": "x"})),
                     ("/v1/chat/completions", dict(json={"prompt": "x"})), ("/chat", dict(content=b""))]:
        a, b = both("POST", path, **kw)
        assert (a.status_code, a.json()) == (b.status_code, b.json()), (path, kw)
    # no GET routes at all (no /health, no /metrics): 501 from both
    for path in ("/health", "/metrics", "/chat"):
        a, b = both("GET", path)
        assert a.status_code == b.status_code == 501, path
    # success: every alias route, both request spellings; output = prompt + completion, nothing else in the body
    for path in ("/chat", "/generate", "/completion"):
        for body in ({"prompt": "Hello", "max_tokens": 3}, {"input": "Hello", "max_new_tokens": "3"},
                     {"prompt": "Hello", "max_tokens": "junk"}, {"prompt": "Hello"}):
            a, b = both("POST", path, json=body)
            assert a.status_code == b.status_code == 200
            ja, jb = a.json(), b.json()
            assert set(ja) == set(jb) == {"output"}
            assert ja["output"].startswith("Hello") and jb["output"].startswith("Hello")
            want_new = stub.calls[-1][1]["max_new_tokens"]           # what the reference asked its pipeline for
            assert eng.seen[-1]["params"].max_tokens == want_new, body
            assert len(jb["output"]) == len("Hello") + min(want_new, 5)   # the fake engine emits <= 5 tokens
    # sampling as the reference asks its pipeline: temperature 0.7, sampling on, no chat template, BOS + raw text
    assert stub.calls[-1][1]["temperature"] == 0.7 and stub.calls[-1][1]["do_sample"] is True
    p = eng.seen[-1]["params"]
    assert p.temperature == 0.7 and p.top_p == 1.0 and not p.ignore_eos
    from llm import hf_cpu_server as ours

    tok = ours._backend._tokenizer
    assert eng.seen[-1]["ids"][0] == tok.bos_token_id and tok.decode(eng.seen[-1]["ids"][1:]) == "Hello"
    assert ref.LLM_MAX_TOKENS == ours.LLM_MAX_TOKENS == 512 and stub.calls[-1][1]["max_new_tokens"] == 512
