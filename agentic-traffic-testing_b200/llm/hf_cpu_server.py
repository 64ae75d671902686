"""Drop-in for the reference's llm/hf_cpu_server.py, served by the B200 engine.

The reference file is a blocking `http.server` around a HF fp32 CPU pipeline
(hf_cpu_server.py:34-105): POST /chat|/generate|/completion with {"prompt"|"input",
"max_tokens"|"max_new_tokens"} -> {"output": text}; env LLM_MODEL|MODEL_NAME, HOST, PORT,
LLM_MAX_TOKENS. Here the same entry point (`python -m llm.hf_cpu_server`, `run()`) starts the
B200 backend with those env semantics. Its request fields are a subset of /chat's (which already
accepts `max_new_tokens`), and the response is a superset ({"output", "meta"}). Two behaviours of
the original are deliberately NOT reproduced: it echoes the prompt in `output` (a side effect of
HF pipeline defaults) and it serialises requests; unknown POST paths still answer 404.
"""
from __future__ import annotations

import os

from llm import serve_llm

MODEL_NAME = os.environ.get("LLM_MODEL") or os.environ.get("MODEL_NAME", "meta-llama/Llama-3.1-8B-Instruct")
HOST = os.environ.get("HOST", "0.0.0.0")
PORT = int(os.environ.get("PORT", "8000"))


def run() -> None:
    print(f"[*] B200 server (hf_cpu_server entry point) for {MODEL_NAME} on http://{HOST}:{PORT}")
    serve_llm.main(["--model", MODEL_NAME, "--host", HOST, "--port", str(PORT)])


if __name__ == "__main__":
    run()
