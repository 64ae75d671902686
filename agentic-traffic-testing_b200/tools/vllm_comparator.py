"""Comparator, not product code: the library the reference's GPU path wraps (vLLM, `llm/serve_llm.py:346-381`)
timed on the same B200 on the bench's headline workload (BASELINE configs[1]: batch 6, 1024-token prompts with
a 768-token shared prefix, 256 greedy tokens each), with dummy (random) weights of the Llama-3.1-8B shape and
the engine arguments the reference passes (`max_model_len`, `max_num_seqs=12`, `max_num_batched_tokens=8192`,
`gpu_memory_utilization=0.90`, `dtype=bfloat16`, prefix caching on as in vLLM V1's default).

    python agentic-traffic-testing_b200/tools/vllm_comparator.py --steps 5 --warmup 3
    python agentic-traffic-testing_b200/tools/vllm_comparator.py --model 70b --tp 8 --batch 32 --steps 2 --warmup 1

The second form is BASELINE configs[3] (Llama-3.1-70B shape, tensor parallel over the 8 GPUs of the box, fan-out batch
32) with the bursts `bench.py`'s tp70b record uses: vLLM's own fused all-reduce + RMSNorm path next to ours.

Prints one JSON line; exits 0 with {"unavailable": ...} if vLLM cannot be brought up offline. Wall-clock timed
(vLLM owns its streams); the number is reported beside bench.py's, never as bench.py's value.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import tempfile
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent.parent
sys.path.insert(0, str(ROOT))

LLAMA_8B = {
    "architectures": ["LlamaForCausalLM"], "model_type": "llama", "hidden_size": 4096, "intermediate_size": 14336,
    "num_hidden_layers": 32, "num_attention_heads": 32, "num_key_value_heads": 8, "head_dim": 128,
    "vocab_size": 128256, "max_position_embeddings": 131072, "rms_norm_eps": 1e-5, "rope_theta": 500000.0,
    "rope_scaling": {"factor": 8.0, "low_freq_factor": 1.0, "high_freq_factor": 4.0,
                     "original_max_position_embeddings": 8192, "rope_type": "llama3"},
    "hidden_act": "silu", "tie_word_embeddings": False, "attention_bias": False, "mlp_bias": False,
    "bos_token_id": 128000, "eos_token_id": [128001, 128008, 128009], "torch_dtype": "bfloat16",
}


LLAMA_70B = dict(LLAMA_8B, hidden_size=8192, intermediate_size=28672, num_hidden_layers=80, num_attention_heads=64)


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--eager", action="store_true", help="no torch.compile / CUDA graphs in vLLM")
    ap.add_argument("--model", default="8b", choices=["8b", "70b"])
    ap.add_argument("--tp", type=int, default=1)
    ap.add_argument("--batch", type=int, default=0, help="default: bench.py's burst of 6")
    a = ap.parse_args()
    os.environ.setdefault("HF_HUB_OFFLINE", "1")
    os.environ.setdefault("TRANSFORMERS_OFFLINE", "1")
    os.environ.setdefault("VLLM_NO_USAGE_STATS", "1")
    os.environ.setdefault("VLLM_LOGGING_LEVEL", "WARNING")
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("VLLM_HOST_IP", "127.0.0.1")
    from bench import BATCH, OUT, PROMPT, SHARED, burst_prompts  # the same bursts bench.py times

    if a.batch:         # the tp70b record's bursts (bench.py:tp70b_measure): rng seed 4 + step, shared 768 + 256 own
        import numpy as np

        BATCH = a.batch

        def burst_prompts(step):  # noqa: F811
            rng = np.random.default_rng(4 + step)
            shared = rng.integers(0, 128000, SHARED).tolist()
            return [shared + rng.integers(0, 128000, PROMPT - SHARED).tolist() for _ in range(BATCH)]
    cfg = LLAMA_70B if a.model == "70b" else LLAMA_8B

    out = {"impl": "vllm-comparator", "metric": "decode tokens/s", "workload":
           f"fan-out burst batch {BATCH}, prompt {PROMPT} (shared prefix {SHARED}), {OUT} greedy tokens",
           "weights": f"dummy (load_format=dummy), Llama-3.1-{a.model.upper()} shape, bf16", "tensor_parallel": a.tp}
    try:
        import vllm
        from vllm import LLM, SamplingParams

        out["vllm"] = vllm.__version__
        d = tempfile.mkdtemp(prefix="llama8b_cfg_")
        (Path(d) / "config.json").write_text(json.dumps(cfg))
        llm = LLM(model=d, load_format="dummy", skip_tokenizer_init=True, dtype="bfloat16", max_model_len=4096,
                  max_num_seqs=max(12, BATCH), max_num_batched_tokens=8192, gpu_memory_utilization=0.90,
                  tensor_parallel_size=a.tp,
                  enable_prefix_caching=True, enforce_eager=a.eager, seed=0, disable_log_stats=False)
        sp = SamplingParams(temperature=0.0, max_tokens=OUT, ignore_eos=True, detokenize=False)
        times, ttfts = [], []
        for step in range(-a.warmup, a.steps):
            prompts = [{"prompt_token_ids": p} for p in burst_prompts(step)]
            t0 = time.perf_counter()
            res = llm.generate(prompts, sp, use_tqdm=False)
            dt = time.perf_counter() - t0
            n = sum(len(r.outputs[0].token_ids) for r in res)
            assert n == BATCH * OUT, n
            if step >= 0:
                times.append(dt)
                for r in res:
                    m = getattr(r, "metrics", None)
                    lat = getattr(m, "first_token_latency", 0.0) if m is not None else 0.0
                    if lat:
                        ttfts.append(lat)
        # time to first token: per-request engine stats when vLLM exposes them, and in any case the wall time of
        # a fresh burst cut at ONE token (= the time until the LAST of the 6 requests has its first token)
        sp1 = SamplingParams(temperature=0.0, max_tokens=1, ignore_eos=True, detokenize=False)
        first = []
        for k in range(5):
            prompts = [{"prompt_token_ids": p} for p in burst_prompts(500 + k)]
            t0 = time.perf_counter()
            llm.generate(prompts, sp1, use_tqdm=False)
            first.append(time.perf_counter() - t0)
        out["burst_first_token_ms_p50"] = round(1e3 * sorted(first)[len(first) // 2], 2)
        total = sum(times)
        out.update({"value": round(BATCH * OUT * len(times) / total, 1), "unit": "tokens/s", "steps": a.steps,
                    "warmup": a.warmup, "ms_per_step": round(1e3 * total / len(times), 2), "timing": "wall clock",
                    "p50_ttft_ms": round(1e3 * sorted(ttfts)[len(ttfts) // 2], 2) if ttfts else None,
                    "enforce_eager": a.eager})
    except BaseException as exc:  # comparator only: report and move on
        out["unavailable"] = f"{type(exc).__name__}: {str(exc)[:300]}"
    print(json.dumps(out), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
