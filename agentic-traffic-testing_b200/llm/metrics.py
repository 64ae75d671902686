"""Prometheus families of the llm-backend, declared from one table.

Names, types, label sets, bucket bounds and help texts are the scrape contract of the reference
(llm/serve_llm.py:92-167; consumed by infra/monitoring/prometheus.yml:43-47, the Grafana dashboard
and scripts/experiment/scrape_metrics.py), so they are reproduced exactly; the prefix comes from
LLM_METRICS_PREFIX. `-1` in a gauge means "unknown / engine default" as in the reference.
"""
from __future__ import annotations

from typing import Dict, Optional

try:
    from prometheus_client import CONTENT_TYPE_LATEST, Counter, Gauge, Histogram, generate_latest
    HAVE_PROM = True
except ImportError:  # pragma: no cover
    HAVE_PROM = False
    CONTENT_TYPE_LATEST = "text/plain; version=0.0.4; charset=utf-8"
    generate_latest = None  # type: ignore

LATENCY_BUCKETS = (0.5, 1.0, 2.5, 5.0, 10.0, 15.0, 20.0, 30.0, 45.0, 60.0, 90.0, 120.0, 180.0)
BATCH_BUCKETS = (1, 2, 3, 4, 5, 6, 8, 10, 12, 16, 20, 32)
INTERARRIVAL_BUCKETS = (0.01, 0.05, 0.1, 0.25, 0.5, 1.0, 2.0, 5.0, 10.0, 30.0, 60.0)

# (attribute, kind, suffix, help, extra)
SPEC = (
    ("requests", "counter", "requests_total", "Total LLM requests", ("status",)),
    ("latency", "histogram", "request_latency_seconds", "End-to-end LLM request latency", LATENCY_BUCKETS),
    ("queue_wait", "histogram", "queue_wait_seconds", "Time spent waiting in vLLM queue", LATENCY_BUCKETS),
    ("inflight", "gauge", "inflight_requests", "In-flight LLM requests", None),
    ("prompt_tokens", "counter", "prompt_tokens_total", "Total prompt tokens", ()),
    ("completion_tokens", "counter", "completion_tokens_total", "Total completion tokens", ()),
    ("batch_size", "histogram", "batch_size", "Number of requests batched together", BATCH_BUCKETS),
    ("cfg_max_num_seqs", "gauge", "config_max_num_seqs",
     "Configured max_num_seqs (vLLM scheduler concurrency); -1 means default", None),
    ("cfg_max_batched", "gauge", "config_max_num_batched_tokens",
     "Configured max_num_batched_tokens (vLLM scheduler); -1 means default", None),
    ("cfg_gpu_util", "gauge", "config_gpu_memory_utilization",
     "Configured GPU memory utilization target (0-1); -1 means default", None),
    ("cfg_max_tokens", "gauge", "config_max_tokens", "Configured max tokens per generation (LLM_MAX_TOKENS)", None),
    ("kv_blocks", "gauge", "kv_cache_num_gpu_blocks",
     "vLLM KV cache: number of GPU blocks allocated; -1 means unknown", None),
    ("kv_block_size", "gauge", "kv_cache_block_size_tokens", "vLLM KV cache: tokens per block; -1 means unknown", None),
    ("kv_total_tokens", "gauge", "kv_cache_total_tokens",
     "vLLM KV cache: total tokens available in GPU KV cache (num_gpu_blocks * block_size); -1 means unknown", None),
    ("kv_est_conc", "gauge", "kv_cache_est_max_concurrency_at_max_model_len",
     "Estimated max concurrent sequences limited by KV cache at max_model_len; -1 means unknown", None),
    ("computed_conc", "gauge", "computed_max_concurrency",
     "KV-cache-derived max concurrency: num_gpu_blocks * block_size / max_model_len "
     "(matches the 'Maximum concurrency for X tokens' line vLLM logs at startup)", None),
    ("interarrival", "histogram", "interarrival_seconds", "Time between consecutive LLM request arrivals",
     INTERARRIVAL_BUCKETS),
    # additive families (not in the reference): what only an engine-owning backend can know
    ("prefix_hit_tokens", "counter", "prefix_cache_hit_tokens_total",
     "Prompt tokens served from the paged-KV prefix cache", ()),
    ("kv_used_blocks", "gauge", "kv_cache_used_blocks", "KV cache blocks currently referenced", None),
)


class Metrics:
    """All families under one prefix. `enabled` False => every method is a no-op and /metrics is 503."""

    def __init__(self, prefix: str = "llm", enabled: bool = True, include_tokens: bool = True, registry=None):
        self.enabled = bool(enabled and HAVE_PROM)
        self.include_tokens = include_tokens
        self.registry = registry
        self.prefix = prefix
        self.f: Dict[str, object] = {}
        if enabled and not HAVE_PROM:
            print("[llm-metrics] Disabled: prometheus_client not installed. "
                  "Install via requirements.txt to enable /metrics.")
        if not self.enabled:
            return
        kw = {} if registry is None else {"registry": registry}
        for attr, kind, suffix, text, extra in SPEC:
            name = f"{prefix}_{suffix}"
            if kind == "counter":
                # prometheus_client appends _total itself when the name lacks it
                self.f[attr] = Counter(name, text, list(extra), **kw)
            elif kind == "gauge":
                self.f[attr] = Gauge(name, text, **kw)
            else:
                self.f[attr] = Histogram(name, text, buckets=list(extra), **kw)

    def __getattr__(self, item):
        try:
            return self.__dict__["f"][item]
        except KeyError:
            raise AttributeError(item)

    # ---- recording helpers (mirroring _record_metrics / _update_kv_cache_gauges) --------------------
    def record_request(self, status: str, latency_s: float, queue_wait_s: float, prompt_tokens: Optional[int],
                       completion_tokens: Optional[int]) -> None:
        if not self.enabled:
            return
        self.requests.labels(status=status).inc()
        self.latency.observe(latency_s)
        self.queue_wait.observe(queue_wait_s)
        if prompt_tokens is not None:
            self.prompt_tokens.inc(prompt_tokens)
        if completion_tokens is not None:
            self.completion_tokens.inc(completion_tokens)

    def set_kv_capacity(self, num_blocks: Optional[int], block_size: Optional[int], max_model_len: Optional[int],
                        total: Optional[int] = None, est: Optional[int] = None) -> None:
        if not self.enabled:
            return
        if num_blocks is not None and block_size is not None and max_model_len is not None:
            tot = num_blocks * block_size
            self.kv_blocks.set(float(num_blocks))
            self.kv_block_size.set(float(block_size))
            self.kv_total_tokens.set(float(tot))
            if max_model_len > 0:
                self.kv_est_conc.set(float(tot // max_model_len))
                self.computed_conc.set(tot / max_model_len)
        else:
            self.kv_blocks.set(float(num_blocks) if num_blocks is not None else -1.0)
            self.kv_block_size.set(float(block_size) if block_size is not None else -1.0)
            self.kv_total_tokens.set(float(total) if total is not None else -1.0)
            self.kv_est_conc.set(float(est) if est is not None else -1.0)

    def exposition(self) -> bytes:
        return generate_latest() if self.registry is None else generate_latest(self.registry)
