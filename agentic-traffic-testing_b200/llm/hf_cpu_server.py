"""Drop-in for the reference's llm/hf_cpu_server.py, served by the B200 engine.

The reference file (hf_cpu_server.py:34-105) is a blocking `http.server` around a HF fp32 CPU
`pipeline("text-generation")`. Its observable contract, reproduced here route for route:

  POST /chat | /generate | /completion                                                  (:64-66)
      body  {"prompt" | "input": non-empty str, "max_tokens" | "max_new_tokens"?: int-like}  (:76-88)
      200   {"output": <prompt + completion>}     — HF pipelines return the full text    (:90-94)
      400   {"error": "Invalid JSON"} | {"error": "Missing 'prompt' field"}             (:70-79)
  POST anything else -> 404 {"error": "Not found"}                                       (:64-66)
  other methods      -> 501 (BaseHTTPRequestHandler has no do_GET: no /health, no /metrics here)
  sampling           temperature 0.7, do_sample=True, no chat template, stops at EOS     (:91)
  tokenisation       the pipeline's tokenizer adds <|begin_of_text|> in front of the prompt
  env                LLM_MODEL | MODEL_NAME, HOST, PORT, LLM_MAX_TOKENS (512)              (:26-29)
  entry points       `python -m llm.hf_cpu_server`, `run()`                              (:97-110)

What differs, on purpose: requests do not serialise (the engine batches them; the reference's
single-threaded HTTPServer queues them in the kernel's accept backlog), and the arithmetic runs in
bf16 on the GPU instead of fp32 on the host. The reference's CPU arithmetic itself is the parity oracle
and the timed CPU baseline of bench.py (`--impl reference`).
"""
from __future__ import annotations

import json
import os
from typing import Any, Optional

from aiohttp import web

MODEL_NAME = os.environ.get("LLM_MODEL") or os.environ.get("MODEL_NAME", "meta-llama/Llama-3.1-8B-Instruct")
HOST = os.environ.get("HOST", "0.0.0.0")
PORT = int(os.environ.get("PORT", "8000"))
LLM_MAX_TOKENS = int(os.environ.get("LLM_MAX_TOKENS", "512"))
HF_TEMPERATURE = 0.7   # generator(prompt, ..., temperature=0.7, do_sample=True)   (:91)

_backend: Any = None   # a llm.backend.B200Backend (tests inject one built on a fake engine)


def _load_backend():
    """Lazy, like the reference's _load_pipeline (:34-51): the first request pays for the load."""
    global _backend
    if _backend is None:
        from llm.backend import B200Backend

        _backend = B200Backend(MODEL_NAME, None, None, None, None, None)
    return _backend


async def handle_post(request: web.Request) -> web.Response:
    if request.path not in ("/chat", "/generate", "/completion"):
        return web.json_response({"error": "Not found"}, status=404)
    raw = await request.read()
    try:
        data = json.loads(raw.decode("utf-8")) if raw else {}
    except (json.JSONDecodeError, UnicodeDecodeError):
        return web.json_response({"error": "Invalid JSON"}, status=400)
    if not isinstance(data, dict):
        data = {}
    prompt = data.get("prompt") or data.get("input")
    if not isinstance(prompt, str) or not prompt:
        return web.json_response({"error": "Missing 'prompt' field"}, status=400)
    max_new_tokens: Optional[Any] = data.get("max_tokens") or data.get("max_new_tokens")
    if max_new_tokens is not None:
        try:
            max_new_tokens = int(max_new_tokens)
        except (TypeError, ValueError):
            max_new_tokens = LLM_MAX_TOKENS
    else:
        max_new_tokens = LLM_MAX_TOKENS
    backend = _load_backend()
    gen = await backend.generate_ids(None, prompt=prompt, max_tokens=max_new_tokens, temperature=HF_TEMPERATURE,
                                     log_progress=False, ignore_eos=False)
    return web.json_response({"output": prompt + gen.text})


async def handle_other(request: web.Request) -> web.Response:
    # BaseHTTPRequestHandler.send_error(501, "Unsupported method (%r)") for verbs without a do_<VERB>
    return web.Response(status=501, reason=f"Unsupported method ({request.method!r})",
                        text=f"Unsupported method ({request.method!r})")


def create_app() -> web.Application:
    app = web.Application()
    app.router.add_post("/{tail:.*}", handle_post)
    app.router.add_route("*", "/{tail:.*}", handle_other)
    return app


def run() -> None:
    print(f"[*] HF CPU server entry point (B200 engine) for {MODEL_NAME} on http://{HOST}:{PORT}")
    try:
        web.run_app(create_app(), host=HOST, port=PORT, print=None)
    except KeyboardInterrupt:
        print("\n[*] Shutting down HF CPU server.")
    finally:
        if _backend is not None:
            _backend.shutdown()


if __name__ == "__main__":
    run()
