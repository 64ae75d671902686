"""Short, profiler-friendly run of the hot path: one fan-out burst (prefill + a few decode steps) of the
8B model, eager launches (no CUDA graph) so that ncu sees plain kernel launches.

    python agentic-traffic-testing_b200/tools/profile_decode.py [--out-tokens 6] [--model 8b] [--graphs]
"""
from __future__ import annotations

import argparse
import sys
import time
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent.parent
sys.path.insert(0, str(ROOT / "agentic-traffic-testing_b200"))
sys.path.insert(0, str(ROOT))

from atx.config import EngineConfig, ModelSource, model_config_by_name  # noqa: E402
from atx.engine import LLMEngine  # noqa: E402
from atx.scheduler import SamplingParams  # noqa: E402
from bench import burst_prompts  # noqa: E402

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--out-tokens", type=int, default=6)
    ap.add_argument("--model", default="8b")
    ap.add_argument("--graphs", action="store_true")
    a = ap.parse_args()
    eng = LLMEngine(ModelSource(model_config_by_name(a.model), None, 0),
                    EngineConfig(max_model_len=4096, max_num_seqs=12, max_num_batched_tokens=8192,
                                 num_gpu_blocks=4096, enable_cuda_graphs=a.graphs))
    params = SamplingParams(temperature=0.0, max_tokens=a.out_tokens, ignore_eos=True)
    t0 = time.perf_counter()
    for i, p in enumerate(burst_prompts(0)):
        eng.add_request(f"p{i}", p, params)
    while eng.has_work():
        eng.step()
    torch.cuda.synchronize()
    print(f"steps={eng.steps} launches={eng.runner.kernel_launches} wall={time.perf_counter() - t0:.3f}s")
