"""Native OpenAI-style routes on the backend port.

The reference offers this shape only through a separate proxy process
(tools/mcp_universe/openai_proxy.py:67-152) that flattens `messages` into one prompt
(`_flatten_messages` :46-64: "[ROLE]\\ncontent\\n" blocks joined by newlines) and forwards to /chat.
Served natively here with the proxy's response layout, plus real `usage` integers and a real
`finish_reason` instead of the proxy's nulls / constant "stop".

  POST /v1/chat/completions  {model?, messages:[{role, content: str | [{type:"text", text}]}],
                              max_tokens?, temperature?, top_p?, seed?}
  POST /v1/completions       {model?, prompt: str, max_tokens?, temperature?, top_p?, seed?}
  GET  /v1/models

Prompt construction: by default exactly what proxy + /chat produce together (flatten, then the
single-turn chat template), so an agent switched from the proxy to this route sees the same
prompts. With "native_template": true in the body (or LLM_OPENAI_NATIVE_TEMPLATE=1) each message
is rendered with the model's chat template instead.
"""
from __future__ import annotations

import os
import time
import uuid
from typing import Any, Dict, List, Optional, Sequence

from aiohttp import web
from opentelemetry import propagate
from opentelemetry.trace import SpanKind


def flatten_messages(messages: Sequence[Dict[str, Any]]) -> str:
    blocks: List[str] = []
    for m in messages:
        content = m.get("content", "")
        if isinstance(content, list):
            texts = []
            for part in content:
                if isinstance(part, str):
                    texts.append(part)
                elif isinstance(part, dict) and part.get("type") == "text" and isinstance(part.get("text"), str):
                    texts.append(part["text"])
            content = "\n".join(texts)
        elif not isinstance(content, str):
            content = str(content)
        blocks.append("[" + str(m.get("role", "user")).upper() + "]\n" + content + "\n")
    return "\n".join(blocks).strip()


def _plain_messages(messages: Sequence[Dict[str, Any]]) -> List[Dict[str, str]]:
    out = []
    for m in messages:
        c = m.get("content", "")
        if isinstance(c, list):
            c = "\n".join(p if isinstance(p, str) else p.get("text", "") for p in c
                          if isinstance(p, str) or (isinstance(p, dict) and p.get("type") == "text"))
        out.append({"role": str(m.get("role", "user")), "content": c if isinstance(c, str) else str(c)})
    return out


def _num(payload: dict, key: str, cast):
    v = payload.get(key)
    if v is None:
        return None
    try:
        return cast(v)
    except (TypeError, ValueError):
        return None


async def _serve(request: web.Request, kind: str) -> web.Response:
    from llm import serve_llm as S

    if S._backend is None:
        return web.json_response({"error": "Backend not initialized"}, status=503)
    ctx = propagate.extract(dict(request.headers))
    with S._tracer.start_as_current_span("llm.handle_request", context=ctx, kind=SpanKind.SERVER) as span:
        request["t_start"] = time.monotonic()
        S._note_arrival(request["t_start"])
        async with S._Traffic.track() as inflight_now:
            span.set_attribute("app.path", request.path)
            try:
                payload = await request.json()
                assert isinstance(payload, dict)
            except Exception:
                return web.json_response({"error": "Invalid JSON body"}, status=400)
            model = payload.get("model", "local-llm")
            max_tokens = _num(payload, "max_tokens", int)
            if max_tokens is None:
                max_tokens = _num(payload, "max_completion_tokens", int)
            sampling = {k: v for k, v in (("temperature", _num(payload, "temperature", float)),
                                          ("top_p", _num(payload, "top_p", float)),
                                          ("seed", _num(payload, "seed", int))) if v is not None}
            if kind == "chat":
                messages = payload.get("messages")
                if not isinstance(messages, list) or not messages:
                    return web.json_response({"error": "Field 'messages' must be a non-empty list"}, status=400)
                native = bool(payload.get("native_template",
                                          os.environ.get("LLM_OPENAI_NATIVE_TEMPLATE", "").lower() in ("1", "true")))
                original = flatten_messages(messages)
                if native and hasattr(S._backend, "apply_messages_template"):
                    prompt = S._backend.apply_messages_template(_plain_messages(messages))
                else:
                    prompt = S._backend.apply_chat_template(original, None)
                rid = f"chatcmpl-{uuid.uuid4().hex[:24]}"
                templated = True
            else:
                original = payload.get("prompt")
                if isinstance(original, list) and original and all(isinstance(p, str) for p in original):
                    original = original[0]
                if not isinstance(original, str) or not original:
                    return web.json_response({"error": "Field 'prompt' must be a non-empty string"}, status=400)
                prompt, rid, templated = original, f"cmpl-{uuid.uuid4().hex[:24]}", False
            log_id = str(request.headers.get("X-Request-ID") or rid)
            span.set_attribute("app.request_id", log_id)
            gen, latency_s, err, p_tok, c_tok = await S.run_generation(
                request, span, prompt=prompt, original_prompt=original, max_tokens=max_tokens, request_id=log_id,
                templated=templated, inflight_now=inflight_now, sampling=sampling)
            if err is not None:
                return web.json_response({"error": f"Generation failed: {err}"}, status=500)
            finish = "length" if gen.finish_reason == "length" else "stop"
            usage = {"prompt_tokens": p_tok, "completion_tokens": c_tok,
                     "total_tokens": (p_tok + c_tok) if (p_tok is not None and c_tok is not None) else None}
            body: Dict[str, Any] = {"id": rid, "created": int(time.time()), "model": model, "usage": usage}
            if kind == "chat":
                body["object"] = "chat.completion"
                body["choices"] = [{"index": 0, "message": {"role": "assistant", "content": gen.text},
                                    "finish_reason": finish}]
            else:
                body["object"] = "text_completion"
                body["choices"] = [{"index": 0, "text": gen.text, "finish_reason": finish, "logprobs": None}]
            return web.json_response(body)


async def handle_chat_completions(request: web.Request) -> web.Response:
    return await _serve(request, "chat")


async def handle_completions(request: web.Request) -> web.Response:
    return await _serve(request, "text")


async def handle_models(request: web.Request) -> web.Response:
    from llm import serve_llm as S

    name = getattr(S._backend, "_model_name", None) or S.DEFAULT_MODEL_NAME
    return web.json_response({"object": "list", "data": [{"id": name, "object": "model", "owned_by": "local"}]})


def register(app: web.Application) -> None:
    app.router.add_post("/v1/chat/completions", handle_chat_completions)
    app.router.add_post("/v1/completions", handle_completions)
    app.router.add_get("/v1/models", handle_models)
