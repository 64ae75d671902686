"""The whole drop-in stack on the GPU: aiohttp app -> B200Backend -> AsyncEngine thread -> CUDA kernels,
with a tiny random-init model and the byte tokenizer. Checks the /chat and /v1/chat/completions contracts
end to end, concurrent requests (continuous batching), streaming, prefix-cache accounting and /metrics."""
import asyncio
import json

import pytest
import torch
from aiohttp.test_utils import TestClient, TestServer

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def live():
    from atx.config import EngineConfig, resolve_model
    from atx.engine import AsyncEngine, LLMEngine
    from llm import serve_llm as S
    from llm.backend import B200Backend

    ecfg = EngineConfig(max_model_len=1024, max_num_seqs=8, max_num_batched_tokens=256, num_gpu_blocks=256)
    aeng = AsyncEngine(LLMEngine(resolve_model("random:tiny:11"), ecfg))
    backend = B200Backend("random:tiny:11", ecfg.max_model_len, None, ecfg.max_num_seqs, ecfg.max_num_batched_tokens,
                          None, tracer=S._tracer, engine=aeng)
    old = S._backend
    S._backend = backend
    S.export_static_gauges(backend, 0.9)
    loop = asyncio.new_event_loop()
    client = TestClient(TestServer(S.create_app()), loop=loop)
    loop.run_until_complete(client.start_server())
    yield S, client, loop, aeng
    loop.run_until_complete(client.close())
    loop.close()
    aeng.shutdown()
    S._backend = old


def test_chat_runs_on_the_gpu_and_is_deterministic_at_temperature_zero(live):
    S, client, loop, aeng = live

    async def post(body):
        r = await client.post("/v1/completions", json=body)
        return r.status, await r.json()
    body = {"prompt": "The quick brown fox", "max_tokens": 12, "temperature": 0}

    async def both():
        return await asyncio.gather(post(body), post(body))
    (s1, a), (s2, b) = loop.run_until_complete(both())
    assert s1 == s2 == 200
    assert a["choices"][0]["text"] == b["choices"][0]["text"] and len(a["choices"][0]["text"]) > 0
    assert a["usage"] == {"prompt_tokens": 19, "completion_tokens": 12, "total_tokens": 31}
    assert a["choices"][0]["finish_reason"] == "length"
    assert aeng.engine.runner.kernel_launches > 0


def test_concurrent_fanout_burst_shares_the_prefix_and_reports_metrics(live):
    S, client, loop, aeng = live
    shared = "You are one of five workers. Task: summarise the quarterly report for the board. " * 3

    async def one(i):
        r = await client.post("/chat", json={"prompt": shared + f"Worker {i}, go.", "max_tokens": 8},
                              headers={"X-Request-ID": f"w{i}"})
        return r.status, await r.json()

    async def burst():
        return await asyncio.gather(*[one(i) for i in range(5)])
    res = loop.run_until_complete(burst())
    assert all(st == 200 for st, _ in res)
    for i, (_, body) in enumerate(res):
        m = body["meta"]
        assert m["request_id"] == f"w{i}" and m["completion_tokens"] <= 8 and m["prompt_tokens"] > len(shared)
        assert 0 <= m["queue_wait_s"] < 5 and m["total_tokens"] == m["prompt_tokens"] + m["completion_tokens"]
    stats = aeng.stats()
    assert stats["prefix_hits"] > 0 and stats["kv_used_blocks"] == 0

    async def metrics():
        r = await client.get("/metrics")
        return await r.text()
    text = loop.run_until_complete(metrics())
    assert 'llm_requests_total{status="success"}' in text and "llm_prefix_cache_hit_tokens_total" in text
    assert "llm_kv_cache_num_gpu_blocks 256.0" in text
    assert "llm_batch_size_bucket" in text and 'llm_batch_size_bucket{le="1.0"}' in text


def test_streaming_from_the_real_engine(live):
    S, client, loop, aeng = live

    async def go():
        r = await client.post("/chat", json={"prompt": "stream me", "max_tokens": 6, "stream": True})
        return (await r.read()).decode()
    events = [e[6:] for e in loop.run_until_complete(go()).strip().split("\n\n")]
    assert events[-1] == "[DONE]"
    final = json.loads(events[-2])
    assert "".join(json.loads(e)["delta"] for e in events[:-2]) == final["output"]
    assert final["meta"]["completion_tokens"] <= 6


def test_oversize_prompt_is_a_500_not_a_crash(live):
    S, client, loop, aeng = live

    async def go():
        r = await client.post("/chat", json={"prompt": "x" * 5000, "max_tokens": 4, "skip_chat_template": True})
        return r.status, await r.json()
    st, body = loop.run_until_complete(go())
    assert st == 500 and body["error"].startswith("Generation failed:") and "max_model_len" in body["error"]

    async def ok():
        r = await client.post("/chat", json={"prompt": "still alive?", "max_tokens": 2})
        return r.status
    assert loop.run_until_complete(ok()) == 200
