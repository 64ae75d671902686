"""Per-rank compute of a tensor-parallel shard measured on ONE GPU (no collectives): runs the engine with a
model config whose shapes equal one TP rank's shard, to separate compute time from all-reduce time.

    python agentic-traffic-testing_b200/tools/shard_compute.py --tp 8 --batch 32 --ctx 1024 --steps 64
"""
from __future__ import annotations

import argparse
import dataclasses
import json
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent.parent
sys.path.insert(0, str(ROOT / "agentic-traffic-testing_b200"))

from atx.config import EngineConfig, ModelSource, model_config_by_name  # noqa: E402
from atx.engine import LLMEngine  # noqa: E402
from atx.scheduler import SamplingParams  # noqa: E402

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--tp", type=int, default=8)
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--ctx", type=int, default=1024)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--layers", type=int, default=80)
    a = ap.parse_args()
    full = model_config_by_name("70b", a.layers)
    shard = dataclasses.replace(full, num_attention_heads=full.num_attention_heads // a.tp,
                                num_key_value_heads=full.num_key_value_heads // a.tp,
                                intermediate_size=full.intermediate_size // a.tp, vocab_size=full.vocab_size // a.tp // 8 * 8)
    eng = LLMEngine(ModelSource(shard, None, 0),
                    EngineConfig(max_model_len=4096, max_num_seqs=a.batch, max_num_batched_tokens=8192,
                                 num_gpu_blocks=a.batch * (a.ctx + a.steps) // 16 + 64))
    rng = np.random.default_rng(0)
    V = shard.vocab_size
    for rep in range(2):
        prompts = [rng.integers(0, V, a.ctx).tolist() for _ in range(a.batch)]
        eng.runner.step_events = []
        for i, p in enumerate(prompts):
            eng.add_request(f"s{rep}-{i}", p, SamplingParams(temperature=0.0, max_tokens=a.steps, ignore_eos=True))
        while eng.has_work():
            eng.step()
        torch.cuda.synchronize()
        ev = eng.runner.step_events
    dec = [s.elapsed_time(e) for (s, e, nd, npf, ctx) in ev if npf == 0 and nd == a.batch]
    ms = sum(dec) / len(dec)
    byts = shard.weight_bytes_per_decode_step() + shard.kv_bytes_per_token() * a.batch * (a.ctx + a.steps / 2)
    print(json.dumps({"shard_of": f"70b/tp{a.tp}", "layers": a.layers, "batch": a.batch, "decode_step_ms": round(ms, 3),
                      "per_layer_us": round(1e3 * ms / a.layers, 1), "algorithmic_GB": round(byts / 1e9, 2),
                      "GBps": round(byts / ms / 1e6, 1)}))
