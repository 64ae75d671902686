"""Device-resident Llama weights in the layout the kernels stream.

HF state-dict names (transformers LlamaForCausalLM) are fused per layer into
  wqkv  [(Hq + 2 Hkv) * D, H]   = cat(q_proj, k_proj, v_proj)      one GEMM, one pass over x
  wo    [H, Hq * D]
  wgu   [2 I, H]                = gate_proj / up_proj rows INTERLEAVED in groups of 64, so one GEMM
                                  tile holds 64 gate + the 64 matching up features and the
                                  SwiGLU runs in the GEMM epilogue (ATX_GEMM_SWIGLU_BF16)
  wdown [H, I]
all bf16 row-major [out_features, in_features] (the layout torch.nn.Linear / HF store, K-major for
the tcgen05 GEMM, so no transposition ever happens).
"""
from __future__ import annotations

from pathlib import Path
from typing import Dict, List, Optional

import torch

from .config import ModelConfig


def interleave_gate_up(gate: torch.Tensor, up: torch.Tensor) -> torch.Tensor:
    """[I, H] x 2 -> [2I, H]: rows [128j, 128j+64) = gate[64j:64j+64], rows [128j+64, 128j+128) = up[64j:...]."""
    I, H = gate.shape
    if I % 64:
        raise ValueError("intermediate_size must be a multiple of 64")
    return torch.stack((gate.reshape(I // 64, 64, H), up.reshape(I // 64, 64, H)), dim=1).reshape(2 * I, H)


class _DictSource:
    def __init__(self, sd: Dict[str, torch.Tensor]):
        self.sd = sd

    def has(self, name: str) -> bool:
        return name in self.sd

    def shape(self, name: str):
        return tuple(self.sd[name].shape)

    def full(self, name: str) -> torch.Tensor:
        return self.sd[name]

    def rows(self, name: str, lo: int, hi: int) -> torch.Tensor:
        return self.sd[name][lo:hi]

    def cols(self, name: str, lo: int, hi: int) -> torch.Tensor:
        return self.sd[name][:, lo:hi]


class _SafetensorsSource:
    """Name -> shard file from `model.safetensors.index.json` when the checkpoint has one (the sharded hub layout;
    other *.safetensors files in the directory, e.g. Meta's `consolidated.*`, are then never opened), else from
    the key lists of the `model*.safetensors` / `*.safetensors` files. Tensors are read as slices."""

    def __init__(self, path: Path):
        import json

        from safetensors import safe_open

        self._open = safe_open
        self._handles: Dict[Path, object] = {}
        self.where: Dict[str, Path] = {}
        index = path / "model.safetensors.index.json"
        if index.exists():
            for name, fname in json.loads(index.read_text())["weight_map"].items():
                self.where[name] = path / fname
        else:
            files = sorted(path.glob("model*.safetensors")) or sorted(
                f for f in path.glob("*.safetensors") if not f.name.startswith("consolidated"))
            if not files:
                raise FileNotFoundError(f"no *.safetensors under {path}")
            for f in files:
                with safe_open(str(f), framework="pt", device="cpu") as sf:
                    for k in sf.keys():
                        self.where.setdefault(k, f)
        self.bytes_read = 0

    def _file(self, name: str):
        f = self.where[name]
        h = self._handles.get(f)
        if h is None:
            h = self._handles[f] = self._open(str(f), framework="pt", device="cpu").__enter__()
        return h

    def has(self, name: str) -> bool:
        return name in self.where

    def shape(self, name: str):
        return tuple(self._file(name).get_slice(name).get_shape())

    def _count(self, t: torch.Tensor) -> torch.Tensor:
        self.bytes_read += t.numel() * t.element_size()
        return t

    def full(self, name: str) -> torch.Tensor:
        return self._count(self._file(name).get_tensor(name))

    def rows(self, name: str, lo: int, hi: int) -> torch.Tensor:
        return self._count(self._file(name).get_slice(name)[lo:hi])

    def cols(self, name: str, lo: int, hi: int) -> torch.Tensor:
        return self._count(self._file(name).get_slice(name)[:, lo:hi])


class LayerWeights:
    __slots__ = ("ln1", "wqkv", "wo", "ln2", "wgu", "wdown")

    def __init__(self):
        self.ln1 = self.wqkv = self.wo = self.ln2 = self.wgu = self.wdown = None


class ModelWeights:
    def __init__(self, cfg: ModelConfig, device: torch.device):
        self.cfg = cfg
        self.device = device
        self.embed: torch.Tensor = None      # [V, H] bf16
        self.layers: List[LayerWeights] = []
        self.norm: torch.Tensor = None       # [H] bf16
        self.lm_head: torch.Tensor = None    # [V, H] bf16 (aliases embed when tied)
        self.tp_size, self.tp_rank = 1, 0

    def nbytes(self) -> int:
        seen, total = set(), 0
        for t in self._all():
            if t.data_ptr() not in seen:
                seen.add(t.data_ptr())
                total += t.numel() * t.element_size()
        return total

    def _all(self):
        yield self.embed
        yield self.norm
        yield self.lm_head
        for l in self.layers:
            yield from (l.ln1, l.wqkv, l.wo, l.ln2, l.wgu, l.wdown)

    # ---- constructors ---------------------------------------------------------------------------
    @classmethod
    def from_state_dict(cls, cfg: ModelConfig, sd: Dict[str, torch.Tensor], device: torch.device, tp_size: int = 1,
                        tp_rank: int = 0) -> "ModelWeights":
        """`sd` uses HF names; values may be fp32/bf16 on any device. With tp_size > 1 only this rank's
        shard is kept (Megatron layout: q/k/v/gate/up split by output rows = heads / features, o/down by
        input columns, LM head by vocabulary rows; embedding and norms replicated)."""
        return cls._from_source(cfg, _DictSource(sd), device, tp_size, tp_rank)

    @classmethod
    def _from_source(cls, cfg: ModelConfig, src, device: torch.device, tp_size: int, tp_rank: int) -> "ModelWeights":
        """`src` answers full(name) / rows(name, lo, hi) / cols(name, lo, hi) / has(name): a state dict in memory, or
        a checkpoint directory read slice by slice (a rank never touches bytes it does not keep)."""
        w = cls(cfg, device)
        w.tp_size, w.tp_rank = tp_size, tp_rank
        r, n = tp_rank, tp_size

        def dev(t: torch.Tensor) -> torch.Tensor:
            return t.to(device=device, dtype=torch.bfloat16).contiguous()

        def rows(name: str) -> torch.Tensor:   # output-feature (row) shard
            k = src.shape(name)[0] // n
            return src.rows(name, r * k, (r + 1) * k)

        def cols(name: str) -> torch.Tensor:   # input-feature (column) shard
            k = src.shape(name)[1] // n
            return src.cols(name, r * k, (r + 1) * k)

        w.embed = dev(src.full("model.embed_tokens.weight"))
        for i in range(cfg.num_hidden_layers):
            p = f"model.layers.{i}."
            l = LayerWeights()
            l.ln1 = dev(src.full(p + "input_layernorm.weight"))
            l.wqkv = dev(torch.cat([rows(p + "self_attn.q_proj.weight"), rows(p + "self_attn.k_proj.weight"),
                                    rows(p + "self_attn.v_proj.weight")], dim=0))
            l.wo = dev(cols(p + "self_attn.o_proj.weight"))
            l.ln2 = dev(src.full(p + "post_attention_layernorm.weight"))
            l.wgu = dev(interleave_gate_up(rows(p + "mlp.gate_proj.weight"), rows(p + "mlp.up_proj.weight")))
            l.wdown = dev(cols(p + "mlp.down_proj.weight"))
            w.layers.append(l)
        w.norm = dev(src.full("model.norm.weight"))
        tied = cfg.tie_word_embeddings or not src.has("lm_head.weight")
        if tied:
            w.lm_head = w.embed if n == 1 else dev(rows("model.embed_tokens.weight"))
        else:
            w.lm_head = dev(rows("lm_head.weight"))
        w._check()
        return w

    @classmethod
    def random(cls, cfg: ModelConfig, seed: int, device: torch.device, std: float = 0.02, tp_size: int = 1,
               tp_rank: int = 0) -> "ModelWeights":
        """Seeded random init generated ON the device (16 GB for 8B in well under a second). Values
        do not match the CPU-seeded oracle weights; parity tests use from_state_dict instead. With
        tp_size > 1 every rank draws the SAME stream (one full matrix at a time, a transient of at most
        2 I x H fp32) and keeps its Megatron shard, so the ranks hold shards of ONE model, the model
        `random(seed)` builds on a single GPU: replicated tensors agree and TP results can be checked
        against the unsharded engine (70B: 17.4 GB per rank at tp 8)."""
        w = cls(cfg, device)
        w.tp_size, w.tp_rank = tp_size, tp_rank
        n, r = tp_size, tp_rank
        g = torch.Generator(device=device).manual_seed(seed * 1000)
        H, I = cfg.hidden_size, cfg.intermediate_size
        D, q_size = cfg.head_dim, cfg.q_size
        kv_size = (cfg.qkv_size - q_size) // 2

        def mat(rows, k):
            return (torch.randn(rows, k, generator=g, device=device, dtype=torch.float32) * std).to(torch.bfloat16)

        def vec(k):
            return (1.0 + 0.1 * torch.randn(k, generator=g, device=device)).to(torch.bfloat16)

        def rshard(t, lo=0, size=None):   # row shard of the block [lo, lo + size)
            size = t.shape[0] - lo if size is None else size
            k = size // n
            return t[lo + r * k: lo + (r + 1) * k]

        def cshard(t):
            k = t.shape[1] // n
            return t[:, r * k:(r + 1) * k].contiguous()

        w.embed = mat(cfg.vocab_size, H)
        for _ in range(cfg.num_hidden_layers):
            l = LayerWeights()
            l.ln1 = vec(H)
            wqkv, wo = mat(cfg.qkv_size, H), mat(H, q_size)
            l.ln2 = vec(H)
            wgu, wdown = mat(2 * I, H), mat(H, I)
            if n == 1:
                l.wqkv, l.wo, l.wgu, l.wdown = wqkv, wo, wgu, wdown
            else:
                l.wqkv = torch.cat([rshard(wqkv, 0, q_size), rshard(wqkv, q_size, kv_size),
                                    rshard(wqkv, q_size + kv_size, kv_size)], dim=0).contiguous()
                l.wo = cshard(wo)
                l.wgu = rshard(wgu).contiguous()   # interleaved 64-row gate/up groups: a row block IS a feature block
                l.wdown = cshard(wdown)
                del wqkv, wo, wgu, wdown
            w.layers.append(l)
        w.norm = vec(H)
        if cfg.tie_word_embeddings:
            w.lm_head = w.embed if n == 1 else rshard(w.embed).contiguous()
        else:
            head = mat(cfg.vocab_size, H)
            w.lm_head = head if n == 1 else rshard(head).contiguous()
        w._check()
        return w

    @classmethod
    def from_hf_dir(cls, cfg: ModelConfig, path: Path, device: torch.device, tp_size: int = 1,
                    tp_rank: int = 0) -> "ModelWeights":
        """Load an HF checkpoint directory (hub snapshot layout) tensor by tensor: each weight is read from its
        safetensors shard as the slice this rank keeps and moved to the device at once, so host memory holds one
        tensor at a time — a 70B checkpoint under TP=8 costs each rank 1/8 of the bytes, not 8 x 141 GB."""
        return cls._from_source(cfg, _SafetensorsSource(Path(path)), device, tp_size, tp_rank)

    def _check(self) -> None:
        c, n = self.cfg, self.tp_size
        assert self.embed.shape == (c.vocab_size, c.hidden_size)
        assert self.lm_head.shape == (c.vocab_size // n, c.hidden_size)
        assert len(self.layers) == c.num_hidden_layers
        for l in self.layers:
            assert l.wqkv.shape == (c.qkv_size // n, c.hidden_size) and l.wo.shape == (c.hidden_size, c.q_size // n)
            assert l.wgu.shape == (2 * c.intermediate_size // n, c.hidden_size)
            assert l.wdown.shape == (c.hidden_size, c.intermediate_size // n)
        if c.hidden_size % 64 or c.intermediate_size % 64 or c.q_size % 64:
            raise ValueError("hidden / intermediate / q sizes must be multiples of 64 (GEMM K blocking)")
        if c.head_dim not in (64, 128):
            raise ValueError("head_dim must be 64 or 128")


def rope_table(cfg: ModelConfig, max_pos: int, device: torch.device) -> torch.Tensor:
    """[max_pos, head_dim] fp32 cos ++ sin table with Llama-3 frequency scaling, computed on the CPU
    in fp32 exactly as transformers does (modeling_rope_utils.py:550-620, modeling_llama.py:124-135)
    so that RoPE in the kernel is a table lookup of the reference's own values."""
    import math

    D = cfg.head_dim
    inv_freq = 1.0 / (cfg.rope_theta ** (torch.arange(0, D, 2, dtype=torch.int64).to(torch.float32) / D))
    sc = cfg.rope_scaling
    if sc and sc.get("rope_type", "default") == "llama3":
        factor, low, high = sc["factor"], sc["low_freq_factor"], sc["high_freq_factor"]
        old = sc["original_max_position_embeddings"]
        low_wl, high_wl = old / low, old / high
        wl = 2 * math.pi / inv_freq
        scaled = torch.where(wl > low_wl, inv_freq / factor, inv_freq)
        smooth = (old / wl - low) / (high - low)
        smoothed = (1 - smooth) * scaled / factor + smooth * scaled
        medium = ~(wl < high_wl) * ~(wl > low_wl)
        inv_freq = torch.where(medium, smoothed, scaled)
    elif sc and sc.get("rope_type", "default") != "default":
        raise ValueError(f"unsupported rope_type {sc.get('rope_type')!r}")
    freqs = torch.arange(max_pos, dtype=torch.float32)[:, None] * inv_freq[None, :]
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1).contiguous().to(device)
