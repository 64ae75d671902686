#!/usr/bin/env python3
"""One request through the reference's LITERAL CPU server (SURVEY.md section 8d, item "also run through the
actual llm/hf_cpu_server.py HTTP path once with a stub tokenizer"): BASELINE configs[0] — Llama-3.2-1B shape,
random init, fp32, one 128-token prompt, 64 new tokens — first as a direct `model.generate`, then through the
unmodified `HFRequestHandler.do_POST` (/root/reference/llm/hf_cpu_server.py:63-94) on a real localhost socket,
with a real transformers `pipeline("text-generation")` built on that model and an offline byte-level tokenizer
(no tokenizer files exist in this environment). Shows what the HTTP + pipeline layers add to the arithmetic.

Runs only where /root/reference exists (the build container); writes one JSON record:
    python agentic-traffic-testing_b200/tools/ref_hf_cpu_probe.py > profiles/r02_ref_hf_cpu_server_probe.json
"""
from __future__ import annotations

import importlib.util
import json
import os
import sys
import threading
import time
from http.server import HTTPServer
from pathlib import Path

REF = Path("/root/reference")
ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))


def byte_tokenizer():
    """A transformers fast tokenizer over raw bytes (ids 0..255), BOS id 256: 1 ASCII char = 1 token."""
    from tokenizers import Tokenizer, decoders, models, pre_tokenizers
    from tokenizers.pre_tokenizers import ByteLevel
    from transformers import PreTrainedTokenizerFast

    alphabet = sorted(ByteLevel.alphabet())
    vocab = {ch: i for i, ch in enumerate(alphabet)}
    vocab["<|begin_of_text|>"] = 256
    tok = Tokenizer(models.BPE(vocab=vocab, merges=[], unk_token=None))
    tok.pre_tokenizer = pre_tokenizers.ByteLevel(add_prefix_space=False, use_regex=False)
    tok.decoder = decoders.ByteLevel()
    return PreTrainedTokenizerFast(tokenizer_object=tok, bos_token="<|begin_of_text|>", eos_token="<|begin_of_text|>",
                                   pad_token="<|begin_of_text|>")


def main() -> None:
    import torch
    from transformers import pipeline

    import bench

    threads = os.cpu_count() or 1
    model = bench.build_cpu_reference(threads, "1b")
    tok = byte_tokenizer()
    prompt = "".join(chr(97 + (i * 7) % 26) for i in range(128))          # 128 chars = 128 byte tokens
    ids = torch.tensor([tok.encode(prompt)])
    assert ids.shape[1] == 128
    kw = dict(max_new_tokens=64, min_new_tokens=64, pad_token_id=256)
    with torch.no_grad():
        model.generate(ids[:, :16], do_sample=False, max_new_tokens=2, pad_token_id=256)     # warm-up
        direct_greedy, direct_sampled = [], []
        for rep in range(2):                      # a shared VM is noisy: two runs each, the faster one counts
            t0 = time.perf_counter()
            model.generate(ids, do_sample=False, **kw)
            direct_greedy.append(time.perf_counter() - t0)
            torch.manual_seed(rep)
            t0 = time.perf_counter()
            model.generate(ids, do_sample=True, temperature=0.7, **kw)
            direct_sampled.append(time.perf_counter() - t0)
        direct_greedy_s, direct_sampled_s = min(direct_greedy), min(direct_sampled)

    spec = importlib.util.spec_from_file_location("ref_hf_cpu_server", REF / "llm" / "hf_cpu_server.py")
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    model.generation_config.min_new_tokens = 64      # random weights may emit the EOS id: keep 64 tokens, as config 1 says
    model.generation_config.pad_token_id = 256
    ref._PIPELINE = pipeline("text-generation", model=model, tokenizer=tok, device=-1)   # what _load_pipeline builds (:46-51)
    httpd = HTTPServer(("127.0.0.1", 0), ref.HFRequestHandler)
    threading.Thread(target=httpd.serve_forever, daemon=True).start()
    import httpx

    url = f"http://127.0.0.1:{httpd.server_address[1]}/chat"
    http = []
    for rep in range(2):
        torch.manual_seed(rep)
        t0 = time.perf_counter()
        r = httpx.post(url, json={"prompt": prompt, "max_tokens": 64}, timeout=600)
        http.append(time.perf_counter() - t0)
    http_s = min(http)
    body = r.json()
    httpd.shutdown()
    assert r.status_code == 200 and set(body) == {"output"} and body["output"].startswith(prompt)
    print(json.dumps({
        "what": "BASELINE configs[0] through the reference's literal llm/hf_cpu_server.py handler (build container, "
                "not the GPU box)",
        "cores": threads, "model": "Llama-3.2-1B shape, random init, fp32", "prompt_tokens": 128, "new_tokens": 64,
        "direct_generate_greedy_s": round(direct_greedy_s, 3),
        "direct_generate_sampled_t0.7_s": round(direct_sampled_s, 3),
        "http_post_chat_s": round(http_s, 3),
        "all_runs_s": {"direct_greedy": [round(x, 2) for x in direct_greedy],
                       "direct_sampled": [round(x, 2) for x in direct_sampled], "http": [round(x, 2) for x in http]},
        "http_and_pipeline_overhead_s": round(http_s - direct_sampled_s, 3),
        "tokens_per_s_through_http": round(64 / http_s, 2),
        "response_keys": sorted(body), "output_echoes_prompt": True,
    }))


if __name__ == "__main__":
    main()
