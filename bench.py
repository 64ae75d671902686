#!/usr/bin/env python3
"""Headline benchmark: Llama-3.1-8B agent fan-out burst on B200 (BASELINE.json configs[1]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One "step" = one fan-out job: 1 orchestrator + 5 worker requests arriving together (batch 6), each a
1024-token prompt (768-token shared prefix + 256 unique tokens, fresh ids every step so nothing is
cached across steps), 256 greedy output tokens each with EOS ignored -> 1536 output tokens per step.
Prints ONE JSON line (rank 0): decode tokens/s (`value`, device-timed whole-job throughput through the
engine), p50 TTFT, `e2e` (same job through the drop-in HTTP API /chat with host text prompts),
`roofline` of the dominant kernel (the tcgen05 weight-streaming GEMM), the whole-decode-step HBM
fraction, `cpu_baseline` (the reference's CPU arithmetic, HF LlamaForCausalLM fp32, bounded sample).
With N > 1 (torchrun) every rank serves its own replica of the same burst (weak scaling, no
data-path collective); times are max over ranks.
"""
from __future__ import annotations

import argparse
import asyncio
import json
import os
import statistics
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT / "agentic-traffic-testing_b200"))
sys.path.insert(0, str(ROOT))
os.environ.setdefault("LLM_OTEL_EXPORT", "0")
os.environ.setdefault("LLM_IGNORE_EOS", "1")
# NCCL's own log stays on (the driver counts the ranks of each communicator from it) without touching stdout, which
# carries the one JSON line: every process logs to a file of its own and replays the communicator lines on stderr
# when it exits. (Pointing NCCL at /dev/stderr re-opens and truncates whatever file stderr was redirected to.)
if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":   # (the image's default: a banner on stdout)
    os.environ["NCCL_DEBUG"] = "INFO"
os.environ.setdefault("NCCL_DEBUG_SUBSYS", "INIT")
if "NCCL_DEBUG_FILE" not in os.environ and os.environ["NCCL_DEBUG"].upper() not in ("WARN", "VERSION"):
    import atexit
    import glob
    import tempfile

    _nccl_dir = os.environ.setdefault("ATX_NCCL_LOG_DIR", tempfile.mkdtemp(prefix="atx_nccl_"))
    os.environ["NCCL_DEBUG_FILE"] = os.path.join(_nccl_dir, "nccl.%h.%p.log")

    def _replay_nccl_log(pid=os.getpid()):
        if os.getpid() != pid:        # forked / spawned children inherit the hook's environment, not this process's log
            return
        for path in glob.glob(os.path.join(_nccl_dir, f"nccl.*.{pid}.log")):
            try:
                with open(path, errors="replace") as f:
                    for line in f:
                        if "nranks" in line or "Init COMPLETE" in line or "NCCL version" in line:
                            sys.stderr.write(line)
            except OSError:
                pass
        sys.stderr.flush()
    atexit.register(_replay_nccl_log)

METRIC = "Llama-3.1-8B decode tokens/s + p50 TTFT under agent fan-out"
UNIT = "tokens/s"
BATCH, PROMPT, SHARED, OUT = 6, 1024, 768, 256
WORKLOAD_TEXT = ("BASELINE configs[1]: Llama-3.1-8B (32 layers, random init seed 0) on 1xB200 per replica; agent-a + 5 "
                 "agent-b fan-out burst = batch 6, 1024-token prompts (768 shared + 256 unique, fresh ids each step), "
                 "256 greedy tokens each")


def peaks() -> dict:
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return {"hbm_gbs": float(d["hbm_gbs"]), "source": "measured (MEASURED_PEAKS.json)"}
    return {"hbm_gbs": 6650.0, "source": "fallback (B200_PROFILING.md)"}


def burst_prompts(step: int, vocab: int = 128000):
    """Token ids of one fan-out burst: shared prefix + unique suffixes; seed depends on the step."""
    import numpy as np

    rng = np.random.default_rng(2 + 1000 * (step + 1000))
    shared = rng.integers(0, vocab, SHARED).tolist()
    return [shared + rng.integers(0, vocab, PROMPT - SHARED).tolist() for _ in range(BATCH)]


def burst_texts(step: int):
    """Same burst shape as text for the HTTP arm: the offline byte tokenizer maps 1 ASCII char to 1
    token and the server puts <|begin_of_text|> in front of a raw prompt, so a 1023-char prompt is a
    1024-token prompt (the shared prefix: 767 chars + that token = 768)."""
    import numpy as np

    rng = np.random.default_rng(7 + 1000 * (step + 1000))
    alphabet = np.frombuffer(b"abcdefghijklmnopqrstuvwxyz ABCDEFGHIJKLMNOPQRSTUVWXYZ0123456789.,", dtype=np.uint8)
    shared = bytes(rng.choice(alphabet, SHARED - 1)).decode()
    return [shared + bytes(rng.choice(alphabet, PROMPT - SHARED)).decode() for _ in range(BATCH)]


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU every 200 ms while the timed region runs."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self._stop_ev = index, [], set(), None, threading.Event()

    def run(self):
        try:
            import pynvml as nv

            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {getattr(nv, k): k for k in dir(nv) if k.startswith("nvmlClocksThrottleReason")
                     or k.startswith("nvmlClocksEventReason")}
            while not self._stop_ev.is_set():
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                try:
                    bits = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                    for bit, name in names.items():
                        if isinstance(bit, int) and bit and bits & bit and "None" not in name and "All" not in name:
                            self.reasons.add(name.replace("nvmlClocksThrottleReason", "").replace(
                                "nvmlClocksEventReason", ""))
                except Exception:
                    pass
                self._stop_ev.wait(0.2)
        except Exception as exc:  # NVML missing: say so in the record rather than inventing clocks
            self.reasons.add(f"nvml_unavailable:{type(exc).__name__}")

    def finish(self) -> dict:
        self._stop_ev.set()
        self.join(timeout=2)
        rs = sorted({r[0].lower() + r[1:] for r in self.reasons})
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": rs, "samples": len(self.samples)}


def max_over_ranks(dist, values, device):
    """Element-wise MAX of a list of floats over all ranks (identity without a process group)."""
    import torch

    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(x) for x in t]


def gather_over_ranks(dist, items, world):
    """Concatenate a python list from every rank (rank order)."""
    if dist is None:
        return list(items)
    parts = [None] * world
    dist.all_gather_object(parts, list(items))
    return [x for part in parts for x in part]


# ------------------------------------------------------------------------------------------------------
def run_burst(engine, prompts):
    """Submit the burst and step the engine to completion. Returns the finished sequences."""
    from atx.scheduler import SamplingParams

    params = SamplingParams(temperature=0.0, max_tokens=OUT, ignore_eos=True)
    t = time.monotonic()
    seqs = [engine.add_request(f"b{time.monotonic_ns()}-{i}", p, params) for i, p in enumerate(prompts)]
    for s in seqs:
        s.arrival = t
    while engine.has_work():
        engine.step()
    return seqs


async def http_bursts(app, steps, first_step, record, bursts_per_step: int = 1):
    """K steps through POST /chat on a real localhost socket, 6 concurrent clients per burst (and, behind a
    front door over several replicas, `bursts_per_step` bursts arriving together)."""
    import aiohttp
    from aiohttp import web

    runner = web.AppRunner(app)
    await runner.setup()
    site = web.TCPSite(runner, "127.0.0.1", 0)
    await site.start()
    port = site._server.sockets[0].getsockname()[1]
    url = f"http://127.0.0.1:{port}/chat"
    async with aiohttp.ClientSession(timeout=aiohttp.ClientTimeout(total=600)) as sess:
        async def one(text):
            async with sess.post(url, json={"prompt": text, "max_tokens": OUT, "skip_chat_template": True}) as r:
                body = await r.json()
                assert r.status == 200, body
                return body["meta"]
        for k in range(steps):
            texts = [t for j in range(bursts_per_step) for t in burst_texts((first_step + k) * bursts_per_step + j)]
            metas = await asyncio.gather(*[one(t) for t in texts])
            record(metas)
    await runner.cleanup()


def gemm_roofline(engine, T: int, hbm_gbs: float) -> dict:
    """Dominant kernel timed live: every GEMM launch of one decode step (4 per layer + LM head, the
    engine's own weights, so 15 GB are streamed and L2 never helps), CUDA events on the launch stream."""
    import torch

    from atx import _cabi

    r, m = engine.runner, engine.mcfg
    lib, st = r.lib, r.stream
    H, I, V = m.hidden_size, m.intermediate_size, m.vocab_size

    def one_pass():
        s = st.cuda_stream
        for p_qkv, p_o, p_gu, p_down in r.plans:
            lib.atx_gemm_run(p_qkv._h, r.qkv.data_ptr(), m.qkv_size, T, 1, 0, r.flags, s)
            lib.atx_gemm_run(p_o._h, r.h.data_ptr(), H, T, 1, 0, r.flags, s)
            lib.atx_gemm_run(p_gu._h, r.act.data_ptr(), I, T, 3, 1, r.flags, s)
            lib.atx_gemm_run(p_down._h, r.h.data_ptr(), H, T, 1, 0, r.flags, s)
        lib.atx_gemm_run(r.plan_lm._h, r.logits.data_ptr(), V, T, 0, 1, r.flags, s)

    with torch.cuda.stream(st):
        for _ in range(3):
            one_pass()
        st.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        iters = 10
        e0.record(st)
        for _ in range(iters):
            one_pass()
        e1.record(st)
        st.synchronize()
    launches = 4 * m.num_hidden_layers + 1
    ms = e0.elapsed_time(e1) / iters
    act = T * 2 * (m.num_hidden_layers * (2 * H + m.q_size + I) + H)           # bf16 activations read
    outb = T * (m.num_hidden_layers * (4 * m.qkv_size + 8 * H + 2 * I) + 4 * V)  # fp32 / bf16 results written
    byts = m.weight_bytes_per_decode_step() + act + outb
    achieved = byts / ms / 1e6
    traffic = None
    tp = ROOT / "profiles" / "gemm_traffic.json"
    if tp.exists():
        traffic = json.loads(tp.read_text()).get("dram_bytes_per_launch")
    return {"bound": "hbm", "kernel": "gemm_wt_kernel<16> (tcgen05 weight-streaming GEMM, decode T=%d)" % T,
            "achieved": round(achieved, 1), "peak": hbm_gbs, "unit": "GB/s", "frac": round(achieved / hbm_gbs, 4),
            "traffic": traffic, "bytes_per_launch": round(byts / launches), "us_per_launch": round(1e3 * ms / launches, 3),
            "launches_per_decode_step": launches}


_CPU_SHAPES = {
    "8b": dict(hidden_size=4096, intermediate_size=14336, num_hidden_layers=32, num_attention_heads=32,
               num_key_value_heads=8, head_dim=128, tie_word_embeddings=False, factor=8.0),
    "1b": dict(hidden_size=2048, intermediate_size=8192, num_hidden_layers=16, num_attention_heads=32,
               num_key_value_heads=8, head_dim=64, tie_word_embeddings=True, factor=32.0),
}


def build_cpu_reference(threads: int, shape: str = "8b", layers=None):
    """The reference's CPU model (llm/hf_cpu_server.py:39-50): transformers LlamaForCausalLM, fp32, device=cpu,
    Llama-3.1-8B / Llama-3.2-1B shape, cheap random fill (values do not affect CPU speed)."""
    import torch
    from transformers import LlamaConfig, LlamaForCausalLM
    from transformers.models.llama.modeling_llama import LlamaRotaryEmbedding

    torch.set_num_threads(threads)
    sh = dict(_CPU_SHAPES[shape])
    factor = sh.pop("factor")
    if layers:
        sh["num_hidden_layers"] = layers
    cfg = LlamaConfig(vocab_size=128256, rms_norm_eps=1e-5, rope_theta=500000.0, max_position_embeddings=131072,
                      rope_scaling={"rope_type": "llama3", "factor": factor, "low_freq_factor": 1.0,
                                    "high_freq_factor": 4.0, "original_max_position_embeddings": 8192}, **sh)
    with torch.device("meta"):
        model = LlamaForCausalLM(cfg)
    model = model.to_empty(device="cpu").to(torch.float32)
    tile = torch.randn(1 << 20) * 0.02
    for name, p in model.named_parameters():
        flat = p.data.view(-1)
        if p.dim() == 1:
            flat.fill_(1.0)
            continue
        k = tile.numel()
        n = flat.numel() // k * k
        flat[:n].view(-1, k).copy_(tile)
        flat[n:].copy_(tile[: flat.numel() - n])
    if cfg.tie_word_embeddings:
        model.tie_weights()
    model.model.rotary_emb = LlamaRotaryEmbedding(config=cfg, device="cpu")
    model.eval()
    return model


def cpu_generate(model, ids, new_tokens: int):
    """Greedy generation exactly as HF's generate(do_sample=False, min_new_tokens=new_tokens) computes it
    (one prompt pass, then one cached step per token), with prefill and decode timed separately."""
    import torch

    with torch.no_grad():
        t1 = time.perf_counter()
        out = model(ids, use_cache=True)
        past, nxt = out.past_key_values, out.logits[:, -1].argmax(-1, keepdim=True)   # token 1 comes from the prompt pass
        t2 = time.perf_counter()
        for _ in range(new_tokens - 1):
            out = model(nxt, past_key_values=past, use_cache=True)
            past, nxt = out.past_key_values, out.logits[:, -1].argmax(-1, keepdim=True)
        t3 = time.perf_counter()
    return t2 - t1, t3 - t2


# one step of the reference arm = one bounded sample of the burst: same model, same batch of 6, the workload's
# 4:1 prompt:output ratio at 1/32 of its length (the full burst is ~150 s of CPU work per step)
SAMPLE_PROMPT, SAMPLE_OUT = 32, 8


def cpu_burst_sample(model, seed: int, prompt_len: int = SAMPLE_PROMPT, new_tokens: int = SAMPLE_OUT):
    import torch

    g = torch.Generator().manual_seed(2 + seed)
    ids = torch.randint(0, 128000, (BATCH, prompt_len), generator=g)
    return cpu_generate(model, ids, new_tokens)


def sample_text(prompt_len: int, new_tokens: int, k: int, w: int) -> str:
    return (f"transformers.LlamaForCausalLM fp32 on CPU (the class llm/hf_cpu_server.py:39-50 loads, greedy instead of "
            f"its temperature-0.7 sampling: same arithmetic per token), Llama-3.1-8B shape x 32 layers, batch {BATCH}, "
            f"{prompt_len}-token prompts + {new_tokens} tokens each per step (the burst's 4:1 prompt:output ratio at "
            f"1/{PROMPT // prompt_len} length; the full 1024 + 256 burst is minutes of CPU work per step); "
            f"{k} timed step(s) after {w} warm-up")


def cpu_baseline_record(threads: int, samples, k: int, w: int, prompt_len: int, new_tokens: int) -> dict:
    pf = sum(s[0] for s in samples)
    dec = sum(s[1] for s in samples)
    tok = BATCH * new_tokens * len(samples)
    per_step = dec / (len(samples) * max(1, new_tokens - 1))
    pf_per_tok = pf / (len(samples) * BATCH * prompt_len)
    est_full = pf_per_tok * BATCH * PROMPT + per_step * (OUT - 1)
    return {"value": round(tok / (pf + dec), 3), "unit": UNIT, "cores": threads, "kind": "reference",
            "sample": sample_text(prompt_len, new_tokens, k, w),
            "prefill_s_per_step": round(pf / len(samples), 3), "decode_s_per_token_step": round(per_step, 4),
            "decode_only_tokens_per_s": round(BATCH / per_step, 3),
            "extrapolated_full_workload_tokens_per_s": round(BATCH * OUT / est_full, 3),
            "extrapolated_full_workload_ttft_s": round(pf_per_tok * BATCH * PROMPT, 2)}


def cpu_config1(threads: int) -> dict:
    """BASELINE.json configs[0] exactly: Llama-3.2-1B shape, fp32, batch 1, prompt
    randint(0, 128000, (1, 128)) seed 1, 64 greedy tokens (EOS cannot shorten it), prefill and decode timed
    separately; host DRAM GB/s = 4.94 GB of fp32 weights streamed per decoded token."""
    import torch

    model = build_cpu_reference(threads, "1b")
    g = torch.Generator().manual_seed(1)
    ids = torch.randint(0, 128000, (1, 128), generator=g)
    cpu_generate(model, ids[:, :16], 2)                      # page-in / thread-pool warm-up
    pf, dec = cpu_generate(model, ids, 64)
    wbytes = sum(p.numel() for p in model.parameters()) * 4  # tied head counted once: 1.236 G params = 4.94 GB
    return {"workload": "BASELINE configs[0]: Llama-3.2-1B shape random init, fp32 CPU, batch 1, 128-token prompt "
                        "(seed 1), 64 greedy tokens",
            "cores": threads, "prefill_s": round(pf, 3), "decode_s": round(dec, 3),
            "tokens_per_s": round(64 / (pf + dec), 2), "decode_tokens_per_s": round(63 / dec, 2),
            "prefill_tokens_per_s": round(128 / pf, 1), "weight_gb": round(wbytes / 1e9, 3),
            "host_dram_gbs": round(wbytes / 1e9 * 63 / dec, 1)}


def tp70b_measure(dist, rank, world, local, pk, steps, warmup, out_tokens=256, layers=None):
    """BASELINE.json configs[3]: Llama-3.1-70B sharded tensor-parallel over the `world` GPUs of the box (one rank
    per GPU, schedulers in lockstep), agent fan-out batch 32. Returns the record (value = output tokens/s)."""
    import numpy as np
    import torch

    from atx.config import EngineConfig, ModelSource, model_config_by_name
    from atx.engine import LLMEngine
    from atx.scheduler import SamplingParams

    mcfg = model_config_by_name("70b", layers)
    B, P, O = 32, 1024, out_tokens
    ecfg = EngineConfig(max_model_len=4096, max_num_seqs=B, max_num_batched_tokens=8192, gpu_memory_utilization=0.90,
                        device=local, tp_size=world, tp_rank=rank)
    engine = LLMEngine(ModelSource(mcfg, None, 0), ecfg)
    r = engine.runner

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def one(step):
        rng = np.random.default_rng(4 + step)     # identical on every rank: schedulers run in lockstep
        shared = rng.integers(0, 128000, SHARED).tolist()
        prompts = [shared + rng.integers(0, 128000, P - SHARED).tolist() for _ in range(B)]
        params = SamplingParams(temperature=0.0, max_tokens=O, ignore_eos=True)
        r.step_events = []
        seqs = [engine.add_request(f"t{step}-{i}", p, params) for i, p in enumerate(prompts)]
        t = time.monotonic()
        for s_ in seqs:
            s_.arrival = t
        while engine.has_work():
            engine.step()
        ev, r.step_events = r.step_events, None
        return seqs, ev
    for w in range(warmup):
        one(1000 + w)
    barrier()
    l0 = r.kernel_launches
    t0 = time.perf_counter()
    evs, ttfts = [], []
    for k in range(steps):
        seqs, ev = one(k)
        evs += ev
        ttfts += [s.first_token_time - s.arrival for s in seqs]
    barrier()
    wall = max_over_ranks(dist, [time.perf_counter() - t0], "cuda")[0]
    dec = [(a.elapsed_time(b), ctx) for (a, b, nd, npf, ctx) in evs if npf == 0 and nd == B]
    ms = max_over_ranks(dist, [sum(d[0] for d in dec) / len(dec)], "cuda")[0]
    out = {}
    byts = (mcfg.weight_bytes_per_decode_step() + mcfg.kv_bytes_per_token() * (sum(d[1] for d in dec) / len(dec))) / world
    # the step's all-reduce + add + norm calls alone (nothing between them): their share of the step if none of it
    # were hidden; inside the step the GEMM behind each call prefetches its weights meanwhile
    n_ar = 2 * mcfg.num_hidden_layers
    ar_us = r.allreduce_alone_us(B, n_ar)
    ar_us = max_over_ranks(dist, [ar_us or 0.0], "cuda")[0]
    mode = "ll: push protocol, data is the flag (atx_allreduce_add_rmsnorm_ll)" if r.fused_ar else "nccl"
    out.update({"metric": "Llama-3.1-70B tensor-parallel decode tokens/s + p50 TTFT under agent fan-out",
                "value": round(steps * B * O / wall, 1), "unit": UNIT, "steps": steps, "warmup": warmup,
                "ms_per_step": round(1e3 * wall / steps, 1),
                "scaling": "strong", "ttft_p50_ms": round(1e3 * statistics.median(ttfts), 1),
                "decode_step": {"ms": round(ms, 3), "algorithmic_bytes_per_gpu": round(byts),
                                "achieved": round(byts / ms / 1e6, 1), "peak": pk["hbm_gbs"], "unit": "GB/s per GPU",
                                "frac": round(byts / ms / 1e6 / pk["hbm_gbs"], 4),
                                "decode_tokens_per_s": round(B / ms * 1e3, 1),
                                "allreduces_per_step": 2 * mcfg.num_hidden_layers + 1,
                                "allreduce_alone_us_per_call": round(ar_us, 2) if ar_us else None,
                                "allreduce_alone_share_of_step": round(n_ar * ar_us / 1e3 / ms, 4) if ar_us else None,
                                "allreduce_message_bytes": B * mcfg.hidden_size * 4},
                "gpu_launches": r.kernel_launches - l0,
                "tp_allreduce": mode,
                "config": {"workload": f"BASELINE configs[3]: Llama-3.1-70B ({mcfg.num_hidden_layers} layers) tensor-"
                                       f"parallel TP={world} over NVLink, agent fan-out batch {B}, {P}-token prompts "
                                       f"(768 shared), {O} greedy tokens; decode-sized steps: all-reduce + residual add + "
                                       f"RMSNorm in one kernel of ours over peer memory (mode: {mode.split(':')[0]}), "
                                       "replayed as CUDA graphs; prefill-sized steps and the logits: NCCL",
                           "parallelism": f"tp{world}"}})
    engine.runner.close()
    del engine, r
    import gc

    gc.collect()
    torch.cuda.empty_cache()
    return out


def routed_dp_measure(world: int, steps: int, ecfg, model: str = "random:8b:0") -> dict:
    """The 8B replicas behind ONE front door (north_star "one data-parallel replica per GPU behind the
    scheduler"): this process serves POST /chat on one port and routes over `world` engine processes, one per GPU
    (atx/router.py:ReplicaPool — prefix affinity, then least outstanding tokens); each step is `world` fan-out
    bursts arriving together (the per-GPU work of the independent-rank measurement). Wall clock, host text in /
    text out."""
    import contextlib
    import dataclasses
    import io

    from atx.router import ReplicaPool
    from llm import serve_llm as S
    from llm.backend import B200Backend

    t0 = time.perf_counter()
    pool = ReplicaPool(model, dataclasses.replace(ecfg, device=0), world)
    start_s = time.perf_counter() - t0
    S._backend = B200Backend(model, ecfg.max_model_len, None, ecfg.max_num_seqs, ecfg.max_num_batched_tokens,
                             ecfg.gpu_memory_utilization, tracer=S._tracer, engine=pool)
    metas = []
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            asyncio.run(http_bursts(S.create_app(), 3, 30_000, lambda m: None, world))   # graphs, HTTP path warm
            tw = time.perf_counter()
            asyncio.run(http_bursts(S.create_app(), steps, 40_000, metas.append, world))
            wall = time.perf_counter() - tw
        loads = list(pool.router.load)
    finally:
        S._backend = None
        pool.shutdown()
    flat = [m for burst in metas for m in burst]
    assert all(m["completion_tokens"] == OUT and m["prompt_tokens"] == PROMPT - 1 for m in flat)
    return {"value": round(len(flat) * OUT / wall, 1), "unit": UNIT, "ms_per_step": round(1e3 * wall / steps, 2),
            "ttft_p50_ms": round(1e3 * statistics.median(m["queue_wait_s"] for m in flat), 2),
            "requests": len(flat), "replicas": world, "replica_start_s": round(start_s, 1),
            "outstanding_tokens_after": loads,
            "path": f"HTTP POST /chat on 127.0.0.1, ONE aiohttp front door -> ReplicaPool router -> {world} engine "
                    f"processes (one per GPU) over pipes; {world} bursts x 6 requests arrive together per step"}


def routed_trace_measure(world: int, device_ecfg, model: str = "random:8b:0") -> dict:
    """BASELINE configs[2] behind the front door: 4 x `world` synthetic AgentVerse workflows (atx/trace.py: closed
    loop per task, 5-20 ms think time, N <= 5 concurrent executions per stage, seeded) replayed concurrently against
    ReplicaPool(world) — every call is routed (prefix affinity, then least outstanding tokens). One untimed replay
    of other workflows first (graphs, prefix caches start cold for the timed one: fresh ids)."""
    import dataclasses

    from atx import trace as TR
    from atx.config import EngineConfig
    from atx.router import ReplicaPool

    ecfg = dataclasses.replace(EngineConfig(max_model_len=8192, max_num_seqs=32, max_num_batched_tokens=8192,
                                            gpu_memory_utilization=0.90), device=0)
    pool = ReplicaPool(model, ecfg, world)
    try:
        asyncio.run(TR.replay(pool, TR.generate_tasks(2 * world, seed=1003), seed=1003, id_prefix="w-"))
        tasks = TR.generate_tasks(4 * world, seed=3)
        t0 = time.perf_counter()
        recs = asyncio.run(TR.replay(pool, tasks, seed=3, id_prefix="t-"))
        wall = time.perf_counter() - t0
    finally:
        pool.shutdown()
    summ = TR.summarize(recs, wall)
    return {"value": round(summ["tokens_per_s"], 1), "unit": UNIT, "wall_s": round(wall, 2), "tasks": 4 * world,
            "requests": summ["requests"], "ttft_p50_ms": round(summ["ttft_p50_ms"], 2),
            "ttft_p95_ms": round(summ["ttft_p95_ms"], 2), "prompt_tokens": summ["prompt_tokens"],
            "cached_prompt_tokens": summ["cached_prompt_tokens"], "output_tokens": summ["output_tokens"],
            "workload": "BASELINE configs[2]: synthetic AgentVerse workflows (seeded generator, the reference ships no "
                        f"traces), {4 * world} concurrent tasks, closed loop, routed over {world} replicas"}


# ------------------------------------------------------------------------------------------------------
def run_other_workload(args, rank, world, local):
    """BASELINE.json configs[2..4]: opt-in workloads (`--workload trace|long|tp70b`). Each prints one JSON
    line of the same shape as the headline run (value = aggregate output tokens/s)."""
    import torch

    from atx.config import EngineConfig, ModelSource, model_config_by_name
    from atx.engine import AsyncEngine, LLMEngine
    from atx.scheduler import SamplingParams

    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pk = peaks()

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    out = {"metric": METRIC, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "higher_is_better": True, "vs_baseline": None, "dtype": "bf16", "data": "synthetic"}
    if args.workload == "trace":
        from atx import trace as TR

        mcfg = model_config_by_name("8b")
        ecfg = EngineConfig(max_model_len=8192, max_num_seqs=32, max_num_batched_tokens=8192,
                            gpu_memory_utilization=0.90, device=local)
        engine = LLMEngine(ModelSource(mcfg, None, 0), ecfg)
        aeng = AsyncEngine(engine)
        tasks_per_gpu = 4

        def one(step):
            tasks = TR.generate_tasks(tasks_per_gpu, seed=3 + 100 * step + rank)
            t0 = time.perf_counter()
            recs = asyncio.run(TR.replay(aeng, tasks, seed=3 + step, id_prefix=f"s{step}r{rank}-"))
            return recs, time.perf_counter() - t0
        for w in range(args.warmup):
            one(-1 - w + 1000)
        barrier()
        l0 = engine.runner.kernel_launches
        t0 = time.perf_counter()
        recs = []
        for k in range(args.steps):
            r_, _ = one(k)
            recs += r_
        barrier()
        wall = max_over_ranks(dist, [time.perf_counter() - t0], "cuda")[0]
        summ = TR.summarize(recs, wall)
        toks = sum(gather_over_ranks(dist, [summ["output_tokens"]], world))
        ttft = gather_over_ranks(dist, [r.ttft_s for r in recs], world)
        aeng.shutdown()
        if rank == 0 and args.trace_out:
            TR.write_reference_formats(recs, Path(args.trace_out))
        out.update({"value": round(toks / wall, 1), "ms_per_step": round(1e3 * wall / args.steps, 2), "scaling": "weak",
                    "ttft_p50_ms": round(1e3 * statistics.median(ttft), 2),
                    "gpu_launches": engine.runner.kernel_launches - l0,
                    "config": {"workload": "BASELINE configs[2]: Llama-3.1-8B data-parallel replicas replaying synthetic "
                                           "AgentVerse workflows (recruitment -> decision rounds -> N<=5 concurrent "
                                           "executions -> evaluation -> final), 4 concurrent tasks per GPU, closed loop, "
                                           "5-20 ms think time, seeded generator atx/trace.py (the reference ships no traces)",
                               "parallelism": f"dp{world}", "requests": summ["requests"] * world,
                               "prompt_tokens": summ["prompt_tokens"], "cached_prompt_tokens": summ["cached_prompt_tokens"]}})
    elif args.workload == "long":
        mcfg = model_config_by_name("8b")
        B, P, O = 16, 32768, args.long_out
        ecfg = EngineConfig(max_model_len=P + O + 16, max_num_seqs=B, max_num_batched_tokens=8192,
                            gpu_memory_utilization=0.92, enable_prefix_caching=False, device=local)
        engine = LLMEngine(ModelSource(mcfg, None, 0), ecfg)
        r = engine.runner
        import numpy as np

        def one(step):
            rng = np.random.default_rng(5 + step)
            prompts = [rng.integers(0, 128000, P).tolist() for _ in range(B)]
            params = SamplingParams(temperature=0.0, max_tokens=O, ignore_eos=True)
            r.step_events = []
            t0 = time.perf_counter()
            seqs = [engine.add_request(f"l{step}-{i}", p, params) for i, p in enumerate(prompts)]
            while engine.has_work():
                engine.step()
            torch.cuda.synchronize()
            ev, r.step_events = r.step_events, None
            return seqs, ev, time.perf_counter() - t0
        for w in range(args.warmup):
            one(1000 + w)
        walls, evs = [], []
        for k in range(args.steps):
            seqs, ev, wall = one(k)
            walls.append(wall)
            evs += ev
        dec = [(a.elapsed_time(b), ctx) for (a, b, nd, npf, ctx) in evs if npf == 0 and nd == B]
        pf = [(a.elapsed_time(b), npf) for (a, b, nd, npf, ctx) in evs if npf > 0]
        ms = sum(d[0] for d in dec) / len(dec)
        byts = mcfg.weight_bytes_per_decode_step() + mcfg.kv_bytes_per_token() * (sum(d[1] for d in dec) / len(dec))
        pf_ms = sum(p[0] for p in pf) / args.steps
        flops = 2 * 7.50e9 * B * P + 2 * P * P * mcfg.hidden_size * mcfg.num_hidden_layers * B
        wall = sum(walls) / len(walls)
        out.update({"value": round(B * O / wall, 2), "ms_per_step": round(1e3 * wall, 1), "scaling": "weak",
                    "ttft_p50_ms": round(1e3 * statistics.median(s.first_token_time - s.arrival for s in seqs), 1),
                    "decode_step": {"ms": round(ms, 3), "algorithmic_bytes": round(byts),
                                    "achieved": round(byts / ms / 1e6, 1), "peak": pk["hbm_gbs"], "unit": "GB/s",
                                    "frac": round(byts / ms / 1e6 / pk["hbm_gbs"], 4),
                                    "decode_tokens_per_s": round(B / ms * 1e3, 1)},
                    "prefill": {"ms": round(pf_ms, 1), "tokens": B * P, "tokens_per_s": round(B * P / pf_ms * 1e3),
                                "tflops": round(flops / pf_ms / 1e9, 1)},
                    "gpu_launches": r.kernel_launches,
                    "config": {"workload": f"BASELINE configs[4]: Llama-3.1-8B long-context, batch {B} x {P}-token prompts "
                                           f"(chunked prefill, 8192-token budget) + {O} greedy tokens over paged KV",
                               "parallelism": "1 GPU", "kv_blocks": r.num_blocks}})
    elif args.workload == "tp70b":
        out.update(tp70b_measure(dist, rank, world, local, pk, args.steps, args.warmup, args.tp_out, args.layers))
        if rank == 0:
            print(json.dumps(out), flush=True)
        if dist is not None:
            dist.destroy_process_group()
        return
    if rank == 0:
        print(json.dumps(out), flush=True)
    engine.runner.close()
    if dist is not None:
        dist.destroy_process_group()


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="8b")
    ap.add_argument("--no-prefix-cache", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--deterministic", action="store_true", help="EngineConfig.deterministic: fixed-order split-K")
    ap.add_argument("--no-multi", action="store_true", help="N > 1: skip the routed-DP and 70B-TP records")
    ap.add_argument("--tp70b-any-n", action="store_true", help="attach the 70B tensor-parallel record at any N > 1 "
                                                               "(default: only at N = 8, BASELINE configs[3])")
    ap.add_argument("--workload", default="fanout", choices=["fanout", "trace", "long", "tp70b"],
                    help="fanout = headline BASELINE configs[1] (default); the others are configs[2..4]")
    ap.add_argument("--trace-out", default="", help="directory for response.json / llm_calls.jsonl of a trace replay")
    ap.add_argument("--long-out", type=int, default=128)
    ap.add_argument("--tp-out", type=int, default=256)
    ap.add_argument("--layers", type=int, default=None, help="tp70b / --impl reference: reduce depth for smoke runs")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    threads = os.cpu_count() or 1

    if args.impl == "reference":
        # The reference's own CPU implementation of the path, timed on this box's host cores. Rank 0 only.
        if rank != 0:
            return
        model = build_cpu_reference(threads, layers=args.layers)   # --layers: reduced depth, smoke runs only
        K, W = max(1, args.steps), max(0, args.warmup)
        pl, nt = SAMPLE_PROMPT, SAMPLE_OUT
        probe = cpu_burst_sample(model, -1, 8, 2)           # page-in + a rate estimate (not counted)
        est = probe[0] * pl / 8 + probe[1] * (nt - 1)
        while est * (K + W) > 240.0 and nt > 1:              # keep the whole run within a few minutes
            pl, nt = pl // 2, nt // 2
            est = probe[0] * pl / 8 + probe[1] * max(1, nt - 1)
        for w in range(W):
            cpu_burst_sample(model, 1000 + w, pl, nt)
        t0 = time.perf_counter()
        samples = [cpu_burst_sample(model, k, pl, nt) for k in range(K)]
        wall = time.perf_counter() - t0
        rec = cpu_baseline_record(threads, samples, K, W, pl, nt)
        rec["value"] = round(BATCH * nt * K / wall, 3)       # wall clock over exactly K steps
        if args.layers:
            rec["sample"] += f" [SMOKE RUN: {args.layers} layers only]"
        del model
        try:
            c1 = cpu_config1(threads)
        except Exception as exc:
            c1 = {"failed": f"{type(exc).__name__}: {exc}"}
        print(json.dumps({
            "impl": "reference", "metric": METRIC, "value": rec["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": K, "warmup": W, "ms_per_step": round(1e3 * wall / K, 1), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD_TEXT, "parallelism": f"cpu x{threads} threads",
                       "sample": rec["sample"], "tokens_per_step": BATCH * nt},
            "cpu_baseline": rec, "config1_cpu": c1,
            "e2e": {"value": rec["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        }))
        return

    if args.workload != "fanout":
        run_other_workload(args, rank, world, local)
        return

    import torch

    from atx.config import EngineConfig, ModelSource, model_config_by_name
    from atx.engine import AsyncEngine, LLMEngine

    torch.cuda.set_device(local)
    dist = None
    host_pg = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        host_pg = dist.new_group(backend="gloo")   # host-side waits that must not occupy a GPU (routed phase)
    mcfg = model_config_by_name(args.model)
    ecfg = EngineConfig(max_model_len=4096, max_num_seqs=12, max_num_batched_tokens=8192,
                        gpu_memory_utilization=0.90, enable_prefix_caching=not args.no_prefix_cache, device=local,
                        deterministic=args.deterministic)
    engine = LLMEngine(ModelSource(mcfg, None, 0), ecfg)
    r = engine.runner
    pk = peaks()

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for w in range(args.warmup):  # also captures the decode CUDA graphs for every KV-split bucket
        run_burst(engine, burst_prompts(-1 - w))
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    l0, h0, d0, s0 = r.kernel_launches, r.h2d_bytes, r.d2h_bytes, engine.steps
    r.step_events = []  # (start, end, n_decode, n_prefill_tokens, ctx_sum) per engine step
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ttfts = []
    e0.record(r.stream)
    t_host0 = time.perf_counter()
    for k in range(args.steps):
        seqs = run_burst(engine, burst_prompts(k))
        ttfts += [s.first_token_time - s.arrival for s in seqs]
        assert all(s.num_generated == OUT for s in seqs)
    e1.record(r.stream)
    barrier()
    host_ms = (time.perf_counter() - t_host0) * 1e3
    clocks = sampler.finish()
    dev_ms = e0.elapsed_time(e1)
    launches = r.kernel_launches - l0
    eng_steps = engine.steps - s0
    step_events, r.step_events = r.step_events, None
    dev_ms, host_ms = max_over_ranks(dist, [dev_ms, host_ms], "cuda")
    ttfts = gather_over_ranks(dist, ttfts, world)
    tokens = args.steps * BATCH * OUT * world
    value = tokens / (dev_ms / 1e3)

    # ---- whole decode step against the HBM roofline (events recorded inside the timed region) ----
    dec = [(s.elapsed_time(e), nd, ctx) for (s, e, nd, npf, ctx) in step_events if npf == 0 and nd == BATCH]
    decode_step = None
    if dec:
        ms = sum(d[0] for d in dec) / len(dec)
        byts = mcfg.weight_bytes_per_decode_step() + mcfg.kv_bytes_per_token() * (sum(d[2] for d in dec) / len(dec))
        decode_step = {"ms": round(ms, 4), "algorithmic_bytes": round(byts), "achieved": round(byts / ms / 1e6, 1),
                       "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": round(byts / ms / 1e6 / pk["hbm_gbs"], 4),
                       "steps_measured": len(dec), "decode_tokens_per_s": round(BATCH / ms * 1e3, 1)}
    pf = [(s.elapsed_time(e), npf) for (s, e, nd, npf, ctx) in step_events if npf > 0]
    prefill = {"ms_per_burst": round(sum(p[0] for p in pf) / args.steps, 3),
               "tokens_computed_per_burst": round(sum(p[1] for p in pf) / args.steps)} if pf else None

    roof = gemm_roofline(engine, BATCH, pk["hbm_gbs"])
    roof["peak_source"] = pk["source"]

    # ---- end to end through the drop-in HTTP API (host text in, text out) -------------------------
    e2e = None
    if not args.no_e2e:
        from llm import serve_llm as S
        from llm.backend import B200Backend

        aeng = AsyncEngine(engine)
        S._backend = B200Backend("random:%s" % args.model, ecfg.max_model_len, None, ecfg.max_num_seqs,
                                 ecfg.max_num_batched_tokens, ecfg.gpu_memory_utilization, tracer=S._tracer,
                                 engine=aeng)
        metas = []
        import contextlib
        import io

        quiet = io.StringIO()
        with contextlib.redirect_stdout(quiet):  # the server logs one line per request event
            asyncio.run(http_bursts(S.create_app(), 1, 10_000, lambda m: None))  # warm the HTTP path
            barrier()
            h0, d0 = r.h2d_bytes, r.d2h_bytes
            tw = time.perf_counter()
            asyncio.run(http_bursts(S.create_app(), args.steps, 20_000, metas.append))
            barrier()
            wall = time.perf_counter() - tw
        aeng.shutdown()
        wall = max_over_ranks(dist, [wall], "cuda")[0]
        flat = [m for burst in metas for m in burst]
        e2e_ttft = gather_over_ranks(dist, [m["queue_wait_s"] for m in flat], world)
        assert all(m["completion_tokens"] == OUT and m["prompt_tokens"] == PROMPT - 1 for m in flat)
        e2e = {"value": round(tokens / wall, 1), "unit": UNIT,
               "h2d_bytes_per_step": round((r.h2d_bytes - h0) / args.steps),
               "d2h_bytes_per_step": round((r.d2h_bytes - d0) / args.steps),
               "ttft_p50_ms": round(1e3 * statistics.median(e2e_ttft), 2),
               "path": "HTTP POST /chat on 127.0.0.1 (aiohttp server + client), text prompts, byte tokenizer"}

    # ---- N > 1: the two multi-GPU rows of SURVEY 8e measured in the same launch ------------------------
    tp70b = routed = None
    pdl_on, kv_blocks = bool(r.flags), r.num_blocks
    if world > 1 and not args.no_multi:
        import gc

        from llm import serve_llm as S_

        S_._backend = None
        r.close()
        aeng = engine = r = None           # noqa: F841  (drop every reference: 16 GB of weights + the KV pool)
        gc.collect()
        torch.cuda.empty_cache()
        barrier()
        if world == 8 or args.tp70b_any_n:
            # configs[3]: the only path with a collective. Same processes, same NCCL communicator.
            try:
                tp70b = tp70b_measure(dist, rank, world, local, pk, max(1, min(args.steps, 3)), 1, 256)
            except Exception as exc:       # every rank fails alike (lockstep); keep the DP line
                tp70b = {"failed": f"{type(exc).__name__}: {exc}"}
            barrier()
        dist.barrier(group=host_pg)
        if rank == 0:
            try:
                routed = routed_dp_measure(world, args.steps, ecfg, "random:%s:0" % args.model)
            except Exception as exc:
                routed = {"failed": f"{type(exc).__name__}: {exc}"}
            try:
                routed["trace_replay"] = routed_trace_measure(world, ecfg, "random:%s:0" % args.model)
            except Exception as exc:
                routed["trace_replay"] = {"failed": f"{type(exc).__name__}: {exc}"}
        dist.barrier(group=host_pg)        # ranks 1..N-1 wait on the host: their GPUs belong to the replicas now
        if routed is not None and "value" in routed:
            # the N-GPU end-to-end number is the one with the front door in the path; the per-rank HTTP arm
            # (every rank its own server and clients) stays beside it
            e2e_indep, e2e = e2e, {
                "value": routed["value"], "unit": UNIT,
                "h2d_bytes_per_step": (e2e or {}).get("h2d_bytes_per_step", 0) * world,
                "d2h_bytes_per_step": (e2e or {}).get("d2h_bytes_per_step", 0) * world,
                "ttft_p50_ms": routed["ttft_p50_ms"], "path": routed["path"],
                "bytes_note": "staging bytes per replica as counted in the independent-rank arm, times the replicas"}
            routed["independent_ranks_e2e"] = e2e_indep

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            model = build_cpu_reference(threads)
            cpu_burst_sample(model, -1, 8, 2)
            samples = [cpu_burst_sample(model, k) for k in range(2)]
            cpu = cpu_baseline_record(threads, samples, 2, 1, SAMPLE_PROMPT, SAMPLE_OUT)
            del model
            cpu["config1"] = cpu_config1(threads)
        except Exception as exc:  # e.g. not enough host RAM for 32 GB of fp32 weights
            cpu = {"value": None, "unit": UNIT, "cores": threads, "kind": "reference",
                   "sample": f"failed: {type(exc).__name__}: {exc}"}

    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(dev_ms / args.steps, 3), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": WORKLOAD_TEXT,
                       "parallelism": f"dp{world} (independent replicas, no collective)",
                       "prefix_cache": not args.no_prefix_cache, "cuda_graphs": True, "pdl": pdl_on,
                       "deterministic": bool(args.deterministic),
                       "l2": "inputs larger than L2: 15 GB of weights streamed per decode step vs 126 MB L2",
                       "max_num_seqs": ecfg.max_num_seqs, "max_num_batched_tokens": ecfg.max_num_batched_tokens,
                       "kv_blocks": kv_blocks},
            "ttft_p50_ms": round(1e3 * statistics.median(ttfts), 2),
            "ttft_max_ms": round(1e3 * max(ttfts), 2),
            "host_ms_per_step": round(host_ms / args.steps, 3),
            "gpu_launches": launches, "engine_steps": eng_steps,
            "clocks": clocks, "roofline": roof, "decode_step": decode_step, "prefill": prefill, "e2e": e2e,
            "cpu_baseline": cpu,
            **({"routed_dp": routed} if routed is not None else {}),
            **({"tp70b": tp70b} if tp70b is not None else {}),
        }), flush=True)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
