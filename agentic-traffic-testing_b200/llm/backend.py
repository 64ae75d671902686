"""B200Backend — what `AsyncVLLMBackend` (llm/serve_llm.py:343-678) is in the reference, with the
same method surface for the HTTP layer, backed by the atx engine instead of vLLM.

Surface kept (SURVEY.md section 8b "in-process backend surface"): generate, count_tokens,
apply_chat_template, _resolve_tokenizer / _tokenizer, _kv_cache_info, effective_max_num_seqs,
effective_max_num_batched_tokens, effective_max_model_len, kv_cache_total_tokens,
kv_cache_est_max_concurrency_at_max_model_len. Differences, all additive: the backend owns its
tokenizer, so token counts are exact ints taken from ids (the reference re-tokenises strings and
gets None on current vLLM); `generate_ids` exposes ids, finish reason and prefix-cache
hits to the OpenAI-style endpoints.
"""
from __future__ import annotations

import os
import time
import uuid
from dataclasses import dataclass
from typing import Any, List, Optional, Sequence, Tuple

LLM_MAX_TOKENS = int(os.environ.get("LLM_MAX_TOKENS", "512"))
LLM_METRICS_INCLUDE_TOKENS = os.environ.get("LLM_METRICS_INCLUDE_TOKENS", "1").lower() in ("1", "true", "yes", "on")
LLM_APPLY_CHAT_TEMPLATE = os.environ.get("LLM_APPLY_CHAT_TEMPLATE", "1").lower() in ("1", "true", "yes", "on")
LLM_DEFAULT_SYSTEM_PROMPT = os.environ.get(
    "LLM_DEFAULT_SYSTEM_PROMPT",
    "You are a helpful AI assistant. Provide clear, concise, and accurate responses.",
)
DEFAULT_TEMPERATURE = 0.2  # SamplingParams(temperature=0.2) in llm/serve_llm.py:379,521


@dataclass
class Generation:
    text: str
    token_ids: List[int]
    prompt_tokens: int
    queue_wait_s: float
    finish_reason: Optional[str]
    cached_prompt_tokens: int = 0


class B200Backend:
    """One model on one GPU (or N data-parallel replicas) behind an asyncio `generate`."""

    def __init__(self, model: str, max_model_len: Optional[int], dtype: Optional[str], max_num_seqs: Optional[int],
                 max_num_batched_tokens: Optional[int], gpu_memory_utilization: Optional[float],
                 tracer: Any = None, engine: Any = None, tokenizer: Any = None):
        self._model_name = model
        self._tracer = tracer
        self._args = dict(max_model_len=max_model_len, dtype=dtype, max_num_seqs=max_num_seqs,
                          max_num_batched_tokens=max_num_batched_tokens,
                          gpu_memory_utilization=gpu_memory_utilization)
        self._tokenizer = tokenizer
        self._tokenizer_ready = tokenizer is not None
        self._model_path = None
        if engine is not None:          # injected (tests / replica pools)
            self._engine = engine
            self._ecfg = getattr(engine, "ecfg", None)
            self._vocab = getattr(engine, "vocab_size", 128256)
            return
        # ---- real engine: needs CUDA + libatx_b200.so; there is no CPU fallback ----------------
        from atx.config import EngineConfig, resolve_model
        from atx.engine import AsyncEngine, LLMEngine

        if dtype and dtype.lower() not in ("auto", "bfloat16", "bf16"):
            print(f"[llm] note: --dtype {dtype} requested; this backend computes in bfloat16 "
                  "(fp32 accumulate / residual)", flush=True)
        src = resolve_model(model)
        self._model_path = src.path
        dp = int(os.environ.get("LLM_DP_SIZE", "1") or "1")
        ecfg = EngineConfig(
            max_model_len=int(max_model_len or min(src.config.max_position_embeddings, 131072)),
            max_num_seqs=int(max_num_seqs or 256),
            max_num_batched_tokens=int(max_num_batched_tokens or 8192),
            gpu_memory_utilization=float(gpu_memory_utilization or 0.92),
            block_size=int(os.environ.get("LLM_KV_BLOCK_SIZE") or os.environ.get("VLLM_BLOCK_SIZE") or 16),
            enable_prefix_caching=os.environ.get("LLM_PREFIX_CACHING", "1").lower() in ("1", "true", "yes", "on"),
            deterministic=os.environ.get("LLM_DETERMINISTIC", "0").lower() in ("1", "true", "yes", "on"),
        )
        self._ecfg = ecfg
        self._vocab = src.config.vocab_size
        tp = int(os.environ.get("LLM_TP_SIZE") or os.environ.get("LLM_TENSOR_PARALLEL_SIZE") or "1")
        if tp > 1 and dp > 1:
            raise ValueError("LLM_TP_SIZE and LLM_DP_SIZE cannot both be > 1")
        if tp > 1:     # 70B-class models: one model sharded over `tp` GPUs of the box
            from atx.tp_group import TPGroup

            group = TPGroup(model, ecfg, tp)
            self._engine = AsyncEngine(group.engine, tp_group=group)
        elif dp > 1:
            from atx.router import ReplicaPool

            self._engine = ReplicaPool(model, ecfg, dp)
        else:
            self._engine = AsyncEngine(LLMEngine(src, ecfg))

    # ---- capacity / config facts exported as gauges (llm/serve_llm.py:383-502) ----------------------
    def effective_max_num_seqs(self) -> Optional[int]:
        return getattr(self._ecfg, "max_num_seqs", self._args["max_num_seqs"])

    def effective_max_num_batched_tokens(self) -> Optional[int]:
        return getattr(self._ecfg, "max_num_batched_tokens", self._args["max_num_batched_tokens"])

    def effective_max_model_len(self) -> Optional[int]:
        return getattr(self._ecfg, "max_model_len", self._args["max_model_len"])

    def _kv_cache_info(self) -> Tuple[Optional[int], Optional[int]]:
        try:
            nb, bs = self._engine.kv_cache_info()
            return int(nb), int(bs)
        except Exception:
            return None, None

    def kv_cache_total_tokens(self) -> Optional[int]:
        nb, bs = self._kv_cache_info()
        if not nb or not bs or nb <= 0 or bs <= 0:
            return None
        return nb * bs

    def kv_cache_est_max_concurrency_at_max_model_len(self) -> Optional[int]:
        total, max_len = self.kv_cache_total_tokens(), self.effective_max_model_len()
        if total is None or not max_len or max_len <= 0:
            return None
        return total // max_len

    # ---- tokenizer ------------------------------------------------------------------------------------
    def _resolve_tokenizer(self) -> None:
        if self._tokenizer_ready:
            return
        self._tokenizer_ready = True
        from atx.tokenizer import load_tokenizer

        self._tokenizer = load_tokenizer(self._model_path, self._vocab)

    def count_tokens(self, text: str) -> Optional[int]:
        if not LLM_METRICS_INCLUDE_TOKENS:
            return None
        if not text:
            return 0
        self._resolve_tokenizer()
        try:
            return len(self._tokenizer.encode(text, add_special_tokens=False))
        except Exception:
            return None

    def apply_chat_template(self, prompt: str, system_prompt: Optional[str] = None) -> str:
        """Single-turn Llama-3 instruct prompt: the checkpoint's template when its tokenizer has one,
        otherwise the literal layout of llm/serve_llm.py:671-678."""
        if not LLM_APPLY_CHAT_TEMPLATE:
            return prompt
        self._resolve_tokenizer()
        sys_prompt = system_prompt or LLM_DEFAULT_SYSTEM_PROMPT
        messages = ([{"role": "system", "content": sys_prompt}] if sys_prompt else []) + [
            {"role": "user", "content": prompt}]
        return self.apply_messages_template(messages)

    def apply_messages_template(self, messages: Sequence[dict]) -> str:
        self._resolve_tokenizer()
        try:
            return self._tokenizer.apply_chat_template(list(messages), tokenize=False, add_generation_prompt=True)
        except Exception as exc:
            print(f"[llm] Warning: apply_chat_template failed: {exc}, using fallback")
            from atx.tokenizer import llama3_chat_string

            return llama3_chat_string(messages, True)

    # ---- generation -----------------------------------------------------------------------------------
    async def generate(self, prompt: str, request_id: Optional[str] = None, log_progress: bool = True,
                       max_tokens: Optional[int] = None) -> Tuple[str, float]:
        """Reference signature: text in, (text, seconds until the first token) out."""
        g = await self.generate_ids(None, prompt=prompt, request_id=request_id, log_progress=log_progress,
                                    max_tokens=max_tokens)
        return g.text, g.queue_wait_s

    async def generate_ids(self, prompt_ids: Optional[Sequence[int]], *, prompt: Optional[str] = None,
                           request_id: Optional[str] = None, log_progress: bool = True,
                           max_tokens: Optional[int] = None, temperature: Optional[float] = None,
                           top_p: Optional[float] = None, seed: Optional[int] = None,
                           ignore_eos: Optional[bool] = None, on_text=None) -> Generation:
        """`on_text(delta: str)` (optional, may be a coroutine function) receives the newly decoded text
        after every engine step: the hook behind `"stream": true`."""
        from atx.scheduler import SamplingParams

        if request_id is None:
            request_id = str(uuid.uuid4())[:8]
        self._resolve_tokenizer()
        added_bos = 0
        if prompt_ids is None:
            prompt_ids = self._tokenizer.encode(prompt or "", add_special_tokens=False)
        if prompt is not None:
            # A TEXT prompt reaches the model the way vLLM feeds it in the reference (llm/serve_llm.py:527 hands
            # the string to the engine, whose tokenizer adds the special tokens): with <|begin_of_text|> in
            # front. The chat template already starts with it; raw prompts (skip_chat_template,
            # LLM_APPLY_CHAT_TEMPLATE=0, /v1/completions) get it here.
            # `prompt_tokens` keeps counting the text's own tokens, as count_tokens() does (:624-635).
            bos = getattr(self._tokenizer, "bos_token_id", None)
            if bos is not None and (len(prompt_ids) == 0 or prompt_ids[0] != bos):
                prompt_ids = [int(bos)] + list(prompt_ids)
                added_bos = 1
        if ignore_eos is None:
            ignore_eos = os.environ.get("LLM_IGNORE_EOS", "").lower() in ("1", "true", "yes", "on")
        params = SamplingParams(
            temperature=DEFAULT_TEMPERATURE if temperature is None else float(temperature),
            top_p=1.0 if top_p is None else float(top_p),
            max_tokens=int(max_tokens) if max_tokens is not None else LLM_MAX_TOKENS,
            seed=seed, ignore_eos=bool(ignore_eos), stream=on_text is not None)

        t0 = time.monotonic()
        last_log = t0
        queue_wait_s = 0.0
        tr = self._tracer
        wait_span = tr.start_span("llm.time_to_first_token") if tr else None
        gen_span = None
        gen_start = None
        final = None
        sent = 0
        try:
            async for out in self._engine.generate(list(prompt_ids), params, request_id):
                now = time.monotonic()
                if final is None:  # first output == first token
                    queue_wait_s = now - t0
                    if wait_span is not None:
                        wait_span.set_attribute("llm_ttft_seconds", queue_wait_s)
                        wait_span.end()
                        wait_span = None
                    if tr:
                        gen_span = tr.start_span("llm.generate")
                        gen_span.set_attribute("app.request_id", request_id)
                    gen_start = now
                final = out
                if on_text is not None:
                    full = self._tokenizer.decode(list(out.token_ids), skip_special_tokens=True)
                    if full.endswith("\ufffd") and not out.finished:
                        full = full.rstrip("\ufffd")   # hold back an incomplete multi-byte character
                    if len(full) > sent:
                        res = on_text(full[sent:])
                        if hasattr(res, "__await__"):
                            await res
                        sent = len(full)
                if log_progress and now - last_log >= 2.0:
                    n = len(out.token_ids)
                    print(f"[llm] req={request_id} PROGRESS tokens={n} speed={n / max(now - t0, 1e-9):.1f} tok/s",
                          flush=True)
                    last_log = now
            if final is None:
                return Generation("", [], len(prompt_ids) - added_bos, queue_wait_s, None)
            ids = list(final.token_ids)
            if log_progress:
                el = time.monotonic() - t0
                print(f"[llm] req={request_id} GENERATED tokens={len(ids)} time={el:.2f}s "
                      f"speed={len(ids) / max(el, 1e-9):.1f} tok/s", flush=True)
            if gen_span is not None and gen_start is not None:
                gen_span.set_attribute("llm.generate_ms", int((time.monotonic() - gen_start) * 1000))
            text = self._tokenizer.decode(ids, skip_special_tokens=True)
            return Generation(text, ids, len(prompt_ids) - added_bos, queue_wait_s, final.finish_reason,
                              getattr(final, "num_cached_tokens", 0))
        finally:
            if wait_span is not None:
                wait_span.end()
            if gen_span is not None:
                gen_span.end()

    def set_batch_size_observer(self, fn) -> None:
        hook = getattr(self._engine, "set_batch_size_observer", None)
        if hook:
            hook(fn)

    def stats(self) -> dict:
        fn = getattr(self._engine, "stats", None)
        return fn() if fn else {}

    def shutdown(self) -> None:
        fn = getattr(self._engine, "shutdown", None)
        if fn:
            fn()
