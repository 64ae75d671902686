"""Model / engine configuration of the B200 backend.

Mirrors the knobs the reference passes to vLLM's AsyncEngineArgs (llm/serve_llm.py:362-374:
model, max_model_len, dtype, max_num_seqs, max_num_batched_tokens, gpu_memory_utilization) with
vLLM's defaults for the ones the reference leaves unset (block_size 16).
"""
from __future__ import annotations

import json
import os
import re
from dataclasses import dataclass, field
from pathlib import Path
from typing import Optional


@dataclass
class ModelConfig:
    """The LlamaConfig fields the kernels depend on (transformers configuration_llama.py)."""

    name: str = "8b"
    vocab_size: int = 128256
    hidden_size: int = 4096
    intermediate_size: int = 14336
    num_hidden_layers: int = 32
    num_attention_heads: int = 32
    num_key_value_heads: int = 8
    head_dim: int = 128
    rms_norm_eps: float = 1e-5
    rope_theta: float = 500000.0
    rope_scaling: Optional[dict] = field(default_factory=lambda: {
        "rope_type": "llama3", "factor": 8.0, "low_freq_factor": 1.0, "high_freq_factor": 4.0,
        "original_max_position_embeddings": 8192})
    tie_word_embeddings: bool = False
    max_position_embeddings: int = 131072
    eos_token_ids: tuple = (128001, 128008, 128009)  # <|end_of_text|>, <|eom_id|>, <|eot_id|>
    bos_token_id: int = 128000

    @property
    def q_size(self) -> int:
        return self.num_attention_heads * self.head_dim

    @property
    def kv_size(self) -> int:
        return self.num_key_value_heads * self.head_dim

    @property
    def qkv_size(self) -> int:
        return self.q_size + 2 * self.kv_size

    def weight_bytes_per_decode_step(self) -> int:
        """bf16 bytes of all layer matrices + LM head: the algorithmic HBM traffic of one decode step."""
        H, I = self.hidden_size, self.intermediate_size
        per_layer = self.qkv_size * H + H * self.q_size + 2 * I * H + H * I
        return 2 * (per_layer * self.num_hidden_layers + self.vocab_size * H)

    def kv_bytes_per_token(self) -> int:
        return 2 * self.num_hidden_layers * self.kv_size * 2


_SHAPES = {
    "1b": dict(hidden_size=2048, intermediate_size=8192, num_hidden_layers=16, num_attention_heads=32,
               num_key_value_heads=8, head_dim=64, tie_word_embeddings=True, _factor=32.0),
    "3b": dict(hidden_size=3072, intermediate_size=8192, num_hidden_layers=28, num_attention_heads=24,
               num_key_value_heads=8, head_dim=128, tie_word_embeddings=True, _factor=32.0),
    "8b": dict(),
    "70b": dict(hidden_size=8192, intermediate_size=28672, num_hidden_layers=80, num_attention_heads=64,
                num_key_value_heads=8, head_dim=128),
    "tiny": dict(vocab_size=1024, hidden_size=512, intermediate_size=1024, num_hidden_layers=2,
                 num_attention_heads=8, num_key_value_heads=2, head_dim=64, eos_token_ids=(), bos_token_id=0),
    "tiny128": dict(vocab_size=1024, hidden_size=512, intermediate_size=1024, num_hidden_layers=2,
                    num_attention_heads=4, num_key_value_heads=1, head_dim=128, eos_token_ids=(), bos_token_id=0),
}


def model_config_by_name(name: str, num_layers: Optional[int] = None) -> ModelConfig:
    key = name.lower()
    if key not in _SHAPES:
        raise ValueError(f"unknown Llama shape {name!r}; known: {sorted(_SHAPES)}")
    kw = dict(_SHAPES[key])
    factor = kw.pop("_factor", None)
    cfg = ModelConfig(name=key, **kw)
    if factor is not None:
        cfg.rope_scaling = dict(cfg.rope_scaling, factor=factor)
    if num_layers is not None:
        cfg.num_hidden_layers = num_layers
    return cfg


def model_config_from_hf_json(path: Path) -> ModelConfig:
    d = json.loads(Path(path).read_text())
    if d.get("model_type", "llama") != "llama":
        raise ValueError(f"{path}: only model_type 'llama' is supported, got {d.get('model_type')!r}")
    heads = d["num_attention_heads"]
    eos = d.get("eos_token_id", 128009)
    rs = d.get("rope_scaling") or d.get("rope_parameters")
    if rs and "rope_type" not in rs and "type" in rs:
        rs = dict(rs, rope_type=rs["type"])
    return ModelConfig(
        name=str(path.parent.name), vocab_size=d["vocab_size"], hidden_size=d["hidden_size"],
        intermediate_size=d["intermediate_size"], num_hidden_layers=d["num_hidden_layers"],
        num_attention_heads=heads, num_key_value_heads=d.get("num_key_value_heads", heads),
        head_dim=d.get("head_dim") or d["hidden_size"] // heads, rms_norm_eps=d.get("rms_norm_eps", 1e-5),
        rope_theta=d.get("rope_theta", (rs or {}).get("rope_theta", 500000.0)),
        rope_scaling=rs if rs and rs.get("rope_type", "default") != "default" else None,
        tie_word_embeddings=d.get("tie_word_embeddings", False),
        max_position_embeddings=d.get("max_position_embeddings", 131072),
        eos_token_ids=tuple(eos) if isinstance(eos, (list, tuple)) else (eos,),
        bos_token_id=d.get("bos_token_id", 128000),
    )


@dataclass
class ModelSource:
    """Where the weights come from: a local HF directory, or seeded random init of a named shape."""

    config: ModelConfig
    path: Optional[Path] = None      # directory with *.safetensors (+ tokenizer files)
    random_seed: Optional[int] = None


def _hub_snapshot(model_id: str) -> Optional[Path]:
    home = Path(os.environ.get("HF_HOME", Path.home() / ".cache" / "huggingface"))
    base = home / "hub" / ("models--" + model_id.replace("/", "--")) / "snapshots"
    if base.is_dir():
        snaps = sorted(p for p in base.iterdir() if (p / "config.json").exists())
        if snaps:
            return snaps[-1]
    return None


def resolve_model(model: str) -> ModelSource:
    """`--model` semantics of llm/serve_llm.py:1052 plus offline random init.

    * `random:<shape>[:<seed>[:<layers>]]`  (shape in 1b/3b/8b/70b/tiny) — seeded random weights
    * a directory with config.json + safetensors
    * a hub id whose snapshot is in the local HF cache (compose mounts `hf_model_cache`)
    * a hub id that is NOT cached: an error, unless LLM_RANDOM_INIT_SEED is set, in which case the
      shape is inferred from the name (…-1B / -3B / -8B / -70B) and random weights are used.
    """
    if model.startswith("random:"):
        parts = model.split(":")
        seed = int(parts[2]) if len(parts) > 2 and parts[2] else 0
        layers = int(parts[3]) if len(parts) > 3 and parts[3] else None
        return ModelSource(model_config_by_name(parts[1], layers), None, seed)
    p = Path(model)
    if p.is_dir() and (p / "config.json").exists():
        return ModelSource(model_config_from_hf_json(p / "config.json"), p, None)
    snap = _hub_snapshot(model)
    if snap is not None:
        return ModelSource(model_config_from_hf_json(snap / "config.json"), snap, None)
    seed_env = os.environ.get("LLM_RANDOM_INIT_SEED")
    if seed_env not in (None, ""):
        m = re.search(r"(\d+)[bB]\b", model)
        shape = {"1": "1b", "3": "3b", "8": "8b", "70": "70b"}.get(m.group(1) if m else "", None)
        if shape is None:
            raise ValueError(f"cannot infer a Llama shape from model name {model!r}")
        return ModelSource(model_config_by_name(shape), None, int(seed_env))
    raise FileNotFoundError(
        f"model {model!r} is neither a local directory nor in the HF cache (offline). "
        "Mount the checkpoint, or set LLM_RANDOM_INIT_SEED=<int> / use --model random:<shape>:<seed>.")


@dataclass
class EngineConfig:
    max_model_len: int = 4096
    max_num_seqs: int = 256
    max_num_batched_tokens: int = 8192
    gpu_memory_utilization: float = 0.90
    block_size: int = 16
    num_gpu_blocks: Optional[int] = None     # None = size from free HBM
    enable_prefix_caching: bool = True
    enable_cuda_graphs: bool = True
    use_pdl: bool = True
    async_scheduling: bool = True   # launch decode step N+1 before the host has read step N's tokens
    deterministic: bool = False     # split-K sums in a fixed order (ATX_FLAG_ORDERED): same tokens bit for bit on
    #                                 every run of the same batches, a few percent slower per decode step
    tp_allreduce: str = "auto"      # tensor parallel: "auto" = our fused all-reduce + add + norm kernels for steps of up
    #                                 to 256 rows (push form up to 64), NCCL beyond; "nccl" = NCCL for every step
    device: int = 0
    tp_size: int = 1          # tensor-parallel ranks sharing one model (70B); needs torch.distributed NCCL
    tp_rank: int = 0
