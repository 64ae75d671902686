"""B200-native LLM backend server — drop-in for the reference's `python3 -m llm.serve_llm`.

Contract kept byte-for-byte where clients depend on it (reference llm/serve_llm.py):
  routes     POST /chat = /completion = /generate; GET /health = /ready = /live; GET /metrics   (:945-955)
  request    {"prompt"|"input": str, "max_tokens"?, "system_prompt"?, "skip_chat_template"?,
              "request_id"?}; headers traceparent / X-Request-ID                                 (:739-801)
  response   {"output": str, "meta": {request_id, latency_ms, queue_wait_s, prompt_tokens,
              completion_tokens, total_tokens, otel{...}}}                                       (:932-942)
  errors     400 Invalid JSON | Missing 'prompt' field; 500 Generation failed: <exc>;
             503 Backend not initialized | Metrics disabled                                      (:722-780, 891-900)
  metrics    llm/metrics.py (names, types, buckets of :92-167); spans llm.handle_request >
             llm.time_to_first_token, llm.generate under tracer "llm-backend"                    (:540-603, 683, 742)
  CLI / env  --model --host --port --max-model-len --dtype --max-num-seqs
             --max-num-batched-tokens --gpu-memory-utilization, env fallbacks with CLI winning    (:1049-1104)
  log lines  [llm] req=<id> START|PROGRESS|GENERATED|DONE|ERROR|PROMPT_TRUNCATED ...

Additions (none changes the above): `"stream": true` on /chat answers as server-sent events
(`data: {"delta": str}` per engine step, then `data: {"output", "meta"}` and `data: [DONE]`; SURVEY.md
section 8 row f4); `max_new_tokens` alias of the HF server (hf_cpu_server.py:81),
native OpenAI-style /v1/chat/completions, /v1/completions, /v1/models (llm/openai_api.py), exact
token counts in `meta`, the llm_batch_size histogram is actually observed.
The engine beneath is the atx scheduler + sm_100a kernels; there is no vLLM and no CPU fallback.
"""
from __future__ import annotations

import argparse
import asyncio
import json
import os
import time
import uuid
from contextlib import asynccontextmanager
from typing import Any, Dict, List, Optional

from aiohttp import web
from opentelemetry import propagate
from opentelemetry.trace import SpanKind

from llm.backend import LLM_APPLY_CHAT_TEMPLATE, LLM_MAX_TOKENS, B200Backend
from llm.metrics import CONTENT_TYPE_LATEST, Metrics
from llm.tracing import get_tracer


def _flag(name: str, default: str = "") -> bool:
    return os.environ.get(name, default).lower() in ("1", "true", "yes", "on")


DEFAULT_MODEL_NAME = os.environ.get("LLM_MODEL", "meta-llama/Llama-3.1-8B-Instruct")
LOG_LLM_REQUESTS = _flag("LOG_LLM_REQUESTS")
LOG_LLM_MAX_CHARS = int(os.environ.get("LLM_LOG_MAX_CHARS", "500"))
LLM_DTYPE = os.environ.get("LLM_DTYPE")
LLM_MAX_NUM_SEQS = os.environ.get("LLM_MAX_NUM_SEQS")
LLM_MAX_NUM_BATCHED_TOKENS = os.environ.get("LLM_MAX_NUM_BATCHED_TOKENS")
LLM_GPU_MEMORY_UTILIZATION = os.environ.get("LLM_GPU_MEMORY_UTILIZATION")
LLM_MAX_MODEL_LEN = int(os.environ.get("LLM_MAX_MODEL_LEN") or "0")
LLM_PROMPT_SAFETY_MARGIN_TOKENS = int(os.environ.get("LLM_PROMPT_SAFETY_MARGIN_TOKENS", "128"))
LLM_METRICS_ENABLED = _flag("LLM_METRICS_ENABLED", "1")
LLM_METRICS_INCLUDE_TOKENS = _flag("LLM_METRICS_INCLUDE_TOKENS", "1")
LLM_METRICS_PREFIX = os.environ.get("LLM_METRICS_PREFIX", "llm")

METRICS = Metrics(LLM_METRICS_PREFIX, LLM_METRICS_ENABLED, LLM_METRICS_INCLUDE_TOKENS)
_METRICS_READY = METRICS.enabled

_backend: Optional[B200Backend] = None  # set at start-up; tests may inject any object with the surface
_tracer = get_tracer("llm-backend")


class _Traffic:
    """In-flight count and last arrival time; only touched from the event-loop thread."""

    inflight = 0
    last_arrival: Optional[float] = None

    @classmethod
    @asynccontextmanager
    async def track(cls):
        cls.inflight += 1
        if _METRICS_READY:
            METRICS.inflight.inc()
        try:
            yield cls.inflight
        finally:  # every exit path, including 400s and exceptions, gives the slot back
            cls.inflight -= 1
            if _METRICS_READY:
                METRICS.inflight.dec()


def _log_prompt(source: str, prompt: str) -> None:
    if not LOG_LLM_REQUESTS:
        return
    keep = max(LOG_LLM_MAX_CHARS, 0)
    shown = prompt[:keep]
    more = f"... [truncated {len(prompt) - keep} chars]" if keep and len(prompt) > keep else ""
    print(f"[llm-request] source={source} prompt_len={len(prompt)} prompt={shown}{more}")


def _log_request_stats(stats: Dict[str, Any]) -> None:
    if LOG_LLM_REQUESTS:
        line = " ".join(f"{k}={v}" for k, v in stats.items() if v is not None)
        if line:
            print(f"[llm-metrics] {line}")


def _otel_span_metadata(span: Any) -> Dict[str, Any]:
    """ids + attributes of the request span, echoed in `meta.otel` (reference :690-712)."""
    meta: Dict[str, Any] = {}
    try:
        ctx = span.get_span_context()
        meta = {"trace_id": f"{int(ctx.trace_id):032x}", "span_id": f"{int(ctx.span_id):016x}",
                "trace_flags": int(getattr(ctx, "trace_flags", 0)), "is_remote": bool(getattr(ctx, "is_remote", False))}
    except Exception:
        pass
    attrs: Dict[str, Any] = {}
    for name in ("attributes", "_attributes"):
        raw = getattr(span, name, None)
        try:
            if raw:
                attrs.update(dict(raw))
        except Exception:
            continue
    if attrs:
        meta["attributes"] = attrs
    return meta


def _coerce_int(value: Any) -> Optional[int]:
    if value is None or isinstance(value, int):
        return value
    try:
        return int(value)
    except (TypeError, ValueError):
        return None


def _truncate_to_context(backend: Any, prompt: str, max_tokens: Optional[int], request_id: str):
    """Keep the HEAD of an over-long prompt: at most LLM_MAX_MODEL_LEN - max_new - margin tokens
    (reference :812-844). Returns (prompt, prompt_ids | None, dropped | None)."""
    ids = None
    if LLM_MAX_MODEL_LEN <= 0:
        return prompt, ids, None
    backend._resolve_tokenizer()
    tok = getattr(backend, "_tokenizer", None)
    if tok is None:
        return prompt, ids, None
    try:
        new_budget = max_tokens if max_tokens is not None else LLM_MAX_TOKENS
        room = max(0, LLM_MAX_MODEL_LEN - new_budget - LLM_PROMPT_SAFETY_MARGIN_TOKENS)
        ids = tok.encode(prompt, add_special_tokens=False)
        if len(ids) <= room:
            return prompt, ids, None
        dropped = len(ids) - room
        ids = ids[:room]
        print(f"[llm] req={request_id} PROMPT_TRUNCATED original_tokens={room + dropped} kept={room} "
              f"dropped={dropped}", flush=True)
        return tok.decode(ids), ids, dropped
    except Exception as exc:
        print(f"[llm] req={request_id} WARNING: prompt truncation failed: {exc}", flush=True)
        return prompt, None, None


async def handle_health(request: web.Request) -> web.Response:
    return web.json_response({"status": "ok"})


async def handle_metrics(request: web.Request) -> web.Response:
    if not _METRICS_READY:
        return web.json_response({"error": "Metrics disabled"}, status=503)
    if _backend is not None:
        try:
            st = _backend.stats()
            if "kv_used_blocks" in st:
                METRICS.kv_used_blocks.set(st["kv_used_blocks"])
        except Exception:
            pass
    # the exposition content type carries a charset parameter, which aiohttp only accepts as a raw header
    return web.Response(body=METRICS.exposition(), headers={"Content-Type": CONTENT_TYPE_LATEST})


async def run_generation(request: web.Request, span: Any, *, prompt: str, original_prompt: str,
                         max_tokens: Optional[int], request_id: str, templated: bool, inflight_now: int,
                         sampling: Optional[dict] = None, on_text=None):
    """Shared tail of /chat and the OpenAI routes: truncation, span attributes, logging, generation,
    metrics. Returns (Generation | None, latency_s, error text | None)."""
    start = request["t_start"]
    prompt, prompt_ids, dropped = _truncate_to_context(_backend, prompt, max_tokens, request_id)
    span.set_attribute("app.prompt_length", len(original_prompt))
    span.set_attribute("app.formatted_prompt_length", len(prompt))
    span.set_attribute("app.chat_template_applied", templated)
    span.set_attribute("app.prompt_truncated", dropped is not None)
    if dropped is not None:
        span.set_attribute("app.prompt_truncated_tokens", int(dropped))
    if LOG_LLM_REQUESTS:
        span.set_attribute("app.prompt_preview", original_prompt[:200])
    _log_prompt("http", original_prompt)
    print(f"[llm] req={request_id} START inflight={inflight_now} prompt_len={len(original_prompt)}"
          f"{' (templated)' if templated else ''}{f' [TRUNCATED -{dropped}tok]' if dropped is not None else ''}",
          flush=True)
    try:
        if hasattr(_backend, "generate_ids"):
            extra = dict(sampling or {})
            if on_text is not None:
                extra["on_text"] = on_text
            gen = await _backend.generate_ids(prompt_ids, prompt=prompt, request_id=request_id, log_progress=True,
                                              max_tokens=max_tokens, **extra)
            p_tok, c_tok = gen.prompt_tokens, len(gen.token_ids)
            if not LLM_METRICS_INCLUDE_TOKENS:
                p_tok = c_tok = None
        else:  # a backend with only the reference's surface
            from llm.backend import Generation

            text, wait = await _backend.generate(prompt, request_id=request_id, log_progress=True,
                                                 max_tokens=max_tokens)
            p_tok, c_tok = _backend.count_tokens(prompt), _backend.count_tokens(text)
            gen = Generation(text, [], p_tok or 0, wait, None)
    except Exception as exc:
        latency_s = time.monotonic() - start
        print(f"[llm] req={request_id} ERROR after {int(latency_s * 1000)}ms: {exc}", flush=True)
        METRICS.record_request("error", latency_s, 0.0, None, None)
        return None, latency_s, str(exc), None, None
    if p_tok is not None:
        span.set_attribute("llm.prompt_tokens", p_tok)
    if c_tok is not None:
        span.set_attribute("llm.completion_tokens", c_tok)
    if p_tok is not None and c_tok is not None:
        span.set_attribute("llm.total_tokens", p_tok + c_tok)
    latency_s = time.monotonic() - start
    print(f"[llm] req={request_id} DONE latency={int(latency_s * 1000)}ms prompt={p_tok} completion={c_tok} "
          f"remaining={_Traffic.inflight - 1}", flush=True)
    METRICS.record_request("success", latency_s, gen.queue_wait_s, p_tok, c_tok)
    if _METRICS_READY and gen.cached_prompt_tokens:
        METRICS.prefix_hit_tokens.inc(gen.cached_prompt_tokens)
    _log_request_stats({"status": "success", "latency_ms": int(latency_s * 1000), "prompt_tokens": p_tok,
                        "completion_tokens": c_tok})
    return gen, latency_s, None, p_tok, c_tok


def _note_arrival(now: float) -> None:
    if _Traffic.last_arrival is not None and _METRICS_READY:
        METRICS.interarrival.observe(now - _Traffic.last_arrival)
    _Traffic.last_arrival = now


async def handle_chat(request: web.Request) -> web.Response:
    """POST /chat | /completion | /generate."""
    if _backend is None:
        return web.json_response({"error": "Backend not initialized"}, status=503)
    ctx = propagate.extract(dict(request.headers))
    with _tracer.start_as_current_span("llm.handle_request", context=ctx, kind=SpanKind.SERVER) as span:
        request["t_start"] = time.monotonic()
        _note_arrival(request["t_start"])
        async with _Traffic.track() as inflight_now:
            span.set_attribute("app.path", request.path)
            try:
                data = await request.json()
            except json.JSONDecodeError:
                return web.json_response({"error": "Invalid JSON"}, status=400)
            if not isinstance(data, dict):
                data = {}
            prompt = data.get("prompt") or data.get("input")
            if not isinstance(prompt, str) or not prompt:
                return web.json_response({"error": "Missing 'prompt' field"}, status=400)
            max_tokens = _coerce_int(data.get("max_tokens"))
            if max_tokens is None:
                max_tokens = _coerce_int(data.get("max_new_tokens"))
            request_id = str(request.headers.get("X-Request-ID") or data.get("request_id") or uuid.uuid4().hex[:8])
            span.set_attribute("app.request_id", request_id)

            skip_template = bool(data.get("skip_chat_template", False))
            formatted = prompt if skip_template else _backend.apply_chat_template(prompt, data.get("system_prompt"))
            sse: Optional[web.StreamResponse] = None
            on_text = None
            if data.get("stream") is True and hasattr(_backend, "generate_ids"):
                sse = web.StreamResponse(headers={"Content-Type": "text/event-stream", "Cache-Control": "no-cache",
                                                  "X-Request-ID": request_id})
                await sse.prepare(request)

                async def on_text(delta: str) -> None:
                    await sse.write(b"data: " + json.dumps({"delta": delta}).encode() + b"\n\n")
            gen, latency_s, err, p_tok, c_tok = await run_generation(
                request, span, prompt=formatted, original_prompt=prompt, max_tokens=max_tokens,
                request_id=request_id, templated=(not skip_template) and LLM_APPLY_CHAT_TEMPLATE,
                inflight_now=inflight_now, on_text=on_text)
            if err is not None:
                if sse is not None:
                    await sse.write(b"data: " + json.dumps({"error": f"Generation failed: {err}"}).encode() + b"\n\n")
                    await sse.write_eof()
                    return sse
                return web.json_response({"error": f"Generation failed: {err}"}, status=500)
            meta = {
                "request_id": request_id,
                "latency_ms": int(latency_s * 1000),
                "queue_wait_s": round(gen.queue_wait_s, 4),
                "prompt_tokens": p_tok,
                "completion_tokens": c_tok,
                "total_tokens": (p_tok + c_tok) if (p_tok is not None and c_tok is not None) else None,
                "otel": _otel_span_metadata(span),
            }
            if sse is not None:
                await sse.write(b"data: " + json.dumps({"output": gen.text, "meta": meta}).encode() + b"\n\n")
                await sse.write(b"data: [DONE]\n\n")
                await sse.write_eof()
                return sse
            return web.json_response({"output": gen.text, "meta": meta})


def create_app() -> web.Application:
    app = web.Application()  # aiohttp defaults on purpose (1 MiB body limit => 413, as the reference)
    for path in ("/health", "/ready", "/live"):
        app.router.add_get(path, handle_health)
    app.router.add_get("/metrics", handle_metrics)
    for path in ("/chat", "/completion", "/generate"):
        app.router.add_post(path, handle_chat)
    from llm import openai_api

    openai_api.register(app)
    return app


def export_static_gauges(backend: Any, gpu_memory_utilization: Optional[float]) -> None:
    """Config + KV-capacity gauges set once at start-up (reference :986-1017)."""
    if not _METRICS_READY:
        return
    seqs, batched = backend.effective_max_num_seqs(), backend.effective_max_num_batched_tokens()
    METRICS.cfg_max_num_seqs.set(float(seqs) if seqs is not None else -1.0)
    METRICS.cfg_max_batched.set(float(batched) if batched is not None else -1.0)
    METRICS.cfg_gpu_util.set(float(gpu_memory_utilization) if gpu_memory_utilization is not None else -1.0)
    METRICS.cfg_max_tokens.set(float(LLM_MAX_TOKENS))
    nb, bs = backend._kv_cache_info()
    max_len = backend.effective_max_model_len()
    METRICS.set_kv_capacity(nb, bs, max_len, backend.kv_cache_total_tokens(),
                            backend.kv_cache_est_max_concurrency_at_max_model_len())
    if nb is not None and bs is not None and max_len:
        print(f"[llm-metrics] KV cache gauges set at startup: num_gpu_blocks={nb} block_size={bs} "
              f"max_model_len={max_len} total_tokens={nb * bs} est_max_concurrency={nb * bs // max_len}")
    backend.set_batch_size_observer(METRICS.batch_size.observe)


async def run_async_server(host: str, port: int, model_name: str, max_model_len: Optional[int], dtype: Optional[str],
                           max_num_seqs: Optional[int], max_num_batched_tokens: Optional[int],
                           gpu_memory_utilization: Optional[float]) -> None:
    global _backend
    print(f"[*] Initializing B200 engine for {model_name}...")
    _backend = B200Backend(model=model_name, max_model_len=max_model_len, dtype=dtype, max_num_seqs=max_num_seqs,
                           max_num_batched_tokens=max_num_batched_tokens,
                           gpu_memory_utilization=gpu_memory_utilization, tracer=_tracer)
    export_static_gauges(_backend, gpu_memory_utilization)
    runner = web.AppRunner(create_app())
    await runner.setup()
    site = web.TCPSite(runner, host, port)
    bar = "=" * 60
    print(f"{bar}\n[*] B200 async backend ready\n    Model: {model_name}\n    URL: http://{host}:{port}\n"
          f"    max_num_seqs: {max_num_seqs or 'default'}\n    max_model_len: {max_model_len or 'default'}\n"
          f"    Batching: ENABLED (continuous batching, chunked prefill, prefix cache)\n{bar}")
    await site.start()
    try:
        while True:
            await asyncio.sleep(3600)
    except asyncio.CancelledError:
        pass
    finally:
        await runner.cleanup()
        _backend.shutdown()


def _env_number(value: Optional[str], cast):
    if value in (None, ""):
        return None
    try:
        return cast(value)
    except ValueError:
        return None


def build_arg_parser() -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser(description="Async LLM backend (B200-native) for agentic traffic testbed")
    ap.add_argument("--model", default=DEFAULT_MODEL_NAME,
                    help="HF model name, local checkpoint dir, or random:<shape>[:seed] (default: %(default)s)")
    ap.add_argument("--host", default="0.0.0.0", help="Bind host (default: %(default)s)")
    ap.add_argument("--port", type=int, default=8000, help="HTTP port (default: %(default)d)")
    ap.add_argument("--max-model-len", type=int, default=None, help="Max sequence length for KV cache sizing.")
    ap.add_argument("--dtype", default=None, help="Accepted for compatibility (float16|bfloat16|auto); runs bf16.")
    ap.add_argument("--max-num-seqs", type=int, default=None, help="Max sequences per iteration.")
    ap.add_argument("--max-num-batched-tokens", type=int, default=None, help="Max batched tokens per iteration.")
    ap.add_argument("--gpu-memory-utilization", type=float, default=None, help="Target HBM fraction (0-1).")
    ap.add_argument("--tensor-parallel-size", type=int, default=None,
                    help="Additive: shard the model over this many GPUs of the box (env LLM_TP_SIZE).")
    ap.add_argument("--data-parallel-size", type=int, default=None,
                    help="Additive: this many full replicas, one per GPU (env LLM_DP_SIZE).")
    return ap


def main(argv: Optional[List[str]] = None) -> None:
    args = build_arg_parser().parse_args(argv)
    # CLI wins over env (reference :1101-1104)
    dtype = args.dtype or LLM_DTYPE
    if args.tensor_parallel_size:
        os.environ["LLM_TP_SIZE"] = str(args.tensor_parallel_size)
    if args.data_parallel_size:
        os.environ["LLM_DP_SIZE"] = str(args.data_parallel_size)
    max_num_seqs = args.max_num_seqs or _env_number(LLM_MAX_NUM_SEQS, int)
    max_batched = args.max_num_batched_tokens or _env_number(LLM_MAX_NUM_BATCHED_TOKENS, int)
    gpu_util = args.gpu_memory_utilization or _env_number(LLM_GPU_MEMORY_UTILIZATION, float)
    try:
        asyncio.run(run_async_server(args.host, args.port, args.model, args.max_model_len, dtype, max_num_seqs,
                                     max_batched, gpu_util))
    except KeyboardInterrupt:
        print("\n[*] Shutting down B200 async backend.")


if __name__ == "__main__":
    main()
