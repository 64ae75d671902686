"""CPU restatement of the reference's Llama arithmetic — TEST INFRASTRUCTURE, not product code.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may import
this package. The product path (agentic-traffic-testing_b200/) never does and fails loudly when the
CUDA library is missing.

What it restates: the arithmetic behind the reference's L5 llm-backend. The reference itself holds
no kernels; `llm/hf_cpu_server.py:39-50,91` runs `transformers.LlamaForCausalLM` in fp32 on the CPU
and `llm/serve_llm.py:374-379,521-531` runs the same model through vLLM. The algorithm therefore
lives in the third-party dependency `transformers` (unpinned by the reference's requirements.txt;
5.5.0 in this image), file `transformers/models/llama/modeling_llama.py` (cited per function below as
`modeling_llama.py:LINE`) and `transformers/modeling_rope_utils.py`.

Pinning: the reference has no tests or golden vectors for this path. This
restatement is pinned against the reference's own arithmetic run in the build container:
`oracle/gen_golden.py` instantiates `LlamaForCausalLM` (the class hf_cpu_server.py loads), feeds it
the weights produced by `init_weights`, and stores its outputs in tests/golden/*.npz;
tests/test_oracle.py asserts this file reproduces them. Everything is plain fp32 torch-on-CPU.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import torch


@dataclass
class LlamaShape:
    """Subset of transformers LlamaConfig that the arithmetic depends on."""

    vocab_size: int = 128256
    hidden_size: int = 4096
    intermediate_size: int = 14336
    num_hidden_layers: int = 32
    num_attention_heads: int = 32
    num_key_value_heads: int = 8
    head_dim: int = 128
    rms_norm_eps: float = 1e-5
    rope_theta: float = 500000.0
    rope_scaling: Optional[dict] = field(
        default_factory=lambda: {
            "rope_type": "llama3",
            "factor": 8.0,
            "low_freq_factor": 1.0,
            "high_freq_factor": 4.0,
            "original_max_position_embeddings": 8192,
        }
    )
    tie_word_embeddings: bool = False
    max_position_embeddings: int = 131072


def shape_by_name(name: str, num_layers: Optional[int] = None) -> LlamaShape:
    """Public Llama shapes named in BASELINE.json (SURVEY.md section 8 header)."""
    if name == "1b":
        s = LlamaShape(hidden_size=2048, intermediate_size=8192, num_hidden_layers=16, num_attention_heads=32,
                       num_key_value_heads=8, head_dim=64, tie_word_embeddings=True)
        s.rope_scaling = dict(s.rope_scaling, factor=32.0)
    elif name == "8b":
        s = LlamaShape()
    elif name == "70b":
        s = LlamaShape(hidden_size=8192, intermediate_size=28672, num_hidden_layers=80, num_attention_heads=64,
                       num_key_value_heads=8, head_dim=128)
    elif name == "tiny":  # fixture shape: every kernel path (GQA 4, head_dim 64, llama3 rope) in < 10 MB
        s = LlamaShape(vocab_size=1024, hidden_size=512, intermediate_size=1024, num_hidden_layers=2,
                       num_attention_heads=8, num_key_value_heads=2, head_dim=64)
    elif name == "tiny128":  # same with head_dim 128
        s = LlamaShape(vocab_size=1024, hidden_size=512, intermediate_size=1024, num_hidden_layers=2,
                       num_attention_heads=4, num_key_value_heads=1, head_dim=128)
    else:
        raise ValueError(f"unknown shape {name!r}")
    if num_layers is not None:
        s.num_hidden_layers = num_layers
    return s


# --------------------------------------------------------------------------------------------------
# weights: seed -> identical tensors on the oracle side and the GPU side
# --------------------------------------------------------------------------------------------------
def weight_names(shape: LlamaShape) -> List[Tuple[str, Tuple[int, ...]]]:
    H, I, D = shape.hidden_size, shape.intermediate_size, shape.head_dim
    Hq, Hkv = shape.num_attention_heads, shape.num_key_value_heads
    names: List[Tuple[str, Tuple[int, ...]]] = [("model.embed_tokens.weight", (shape.vocab_size, H))]
    for i in range(shape.num_hidden_layers):
        p = f"model.layers.{i}."
        names += [
            (p + "input_layernorm.weight", (H,)),
            (p + "self_attn.q_proj.weight", (Hq * D, H)),
            (p + "self_attn.k_proj.weight", (Hkv * D, H)),
            (p + "self_attn.v_proj.weight", (Hkv * D, H)),
            (p + "self_attn.o_proj.weight", (H, Hq * D)),
            (p + "post_attention_layernorm.weight", (H,)),
            (p + "mlp.gate_proj.weight", (I, H)),
            (p + "mlp.up_proj.weight", (I, H)),
            (p + "mlp.down_proj.weight", (H, I)),
        ]
    names.append(("model.norm.weight", (H,)))
    if not shape.tie_word_embeddings:
        names.append(("lm_head.weight", (shape.vocab_size, H)))
    return names


def init_weights(shape: LlamaShape, seed: int, std: float = 0.02, norm_jitter: float = 0.1) -> Dict[str, torch.Tensor]:
    """Seeded random init in HF state-dict naming, already ROUNDED TO bf16 and held in fp32.

    Matrices ~ N(0, std) as LlamaPreTrainedModel._init_weights (initializer_range 0.02). Norm weights
    are 1 + jitter*N(0,1) rather than HF's exact ones so that a kernel that forgot the norm weight
    cannot pass. Both sides consume exactly these values, so weight rounding is not an error source.
    """
    g = torch.Generator().manual_seed(seed)
    out: Dict[str, torch.Tensor] = {}
    for name, shp in weight_names(shape):
        if len(shp) == 1:
            w = 1.0 + norm_jitter * torch.randn(shp, generator=g)
        else:
            w = std * torch.randn(shp, generator=g)
        out[name] = w.to(torch.bfloat16).to(torch.float32)
    return out


# --------------------------------------------------------------------------------------------------
# ops
# --------------------------------------------------------------------------------------------------
def rms_norm(x: torch.Tensor, w: torch.Tensor, eps: float) -> torch.Tensor:
    """modeling_llama.py:62-67 — variance in fp32, x * rsqrt(var + eps), then weight multiply."""
    x = x.to(torch.float32)
    var = x.pow(2).mean(-1, keepdim=True)
    return w * (x * torch.rsqrt(var + eps))


def llama3_inv_freq(shape: LlamaShape) -> torch.Tensor:
    """modeling_rope_utils.py:550-620 (_compute_llama3_parameters); default rope when no scaling."""
    D = shape.head_dim
    inv_freq = 1.0 / (shape.rope_theta ** (torch.arange(0, D, 2, dtype=torch.int64).to(torch.float32) / D))
    sc = shape.rope_scaling
    if not sc or sc.get("rope_type", "default") == "default":
        return inv_freq
    if sc["rope_type"] != "llama3":
        raise ValueError("only default / llama3 rope are in scope")
    factor, low, high = sc["factor"], sc["low_freq_factor"], sc["high_freq_factor"]
    old_len = sc["original_max_position_embeddings"]
    low_wavelen, high_wavelen = old_len / low, old_len / high
    wavelen = 2 * math.pi / inv_freq
    scaled = torch.where(wavelen > low_wavelen, inv_freq / factor, inv_freq)
    smooth = (old_len / wavelen - low) / (high - low)
    smoothed = (1 - smooth) * scaled / factor + smooth * scaled
    is_medium = ~(wavelen < high_wavelen) * ~(wavelen > low_wavelen)
    return torch.where(is_medium, smoothed, scaled)


def rope_cos_sin(positions: torch.Tensor, inv_freq: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """modeling_llama.py:124-135 — freqs = pos x inv_freq in fp32, emb = cat(freqs, freqs)."""
    freqs = positions.to(torch.float32)[:, None] * inv_freq[None, :].to(torch.float32)
    emb = torch.cat((freqs, freqs), dim=-1)
    return emb.cos(), emb.sin()  # attention_factor is 1.0 for llama3


def cos_sin_table(shape: LlamaShape, max_pos: int) -> torch.Tensor:
    """[max_pos, head_dim] fp32: row p = cos(p*inv_freq) ++ sin(p*inv_freq) (the kernel's table)."""
    freqs = torch.arange(max_pos, dtype=torch.float32)[:, None] * llama3_inv_freq(shape)[None, :]
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1).contiguous()


def rotate_half(x: torch.Tensor) -> torch.Tensor:
    """modeling_llama.py:138-143."""
    x1, x2 = x[..., : x.shape[-1] // 2], x[..., x.shape[-1] // 2:]
    return torch.cat((-x2, x1), dim=-1)


def apply_rope(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor) -> torch.Tensor:
    """modeling_llama.py:146-168; x: [T, heads, D], cos/sin: [T, D]."""
    return x * cos[:, None, :] + rotate_half(x) * sin[:, None, :]


def attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, q_offset: int) -> torch.Tensor:
    """modeling_llama.py:199-222 eager_attention_forward + :187-196 repeat_kv, causal.

    q: [Tq, Hq, D]; k, v: [Tk, Hkv, D]; query j sits at absolute position q_offset + j and sees keys
    [0, q_offset + j]. Softmax in fp32. Returns [Tq, Hq, D].
    """
    Tq, Hq, D = q.shape
    Tk, Hkv, _ = k.shape
    rep = Hq // Hkv
    kk = k.repeat_interleave(rep, dim=1)
    vv = v.repeat_interleave(rep, dim=1)
    scores = torch.einsum("qhd,khd->hqk", q, kk) * (D ** -0.5)
    qpos = torch.arange(Tq)[:, None] + q_offset
    mask = torch.arange(Tk)[None, :] > qpos
    scores = scores.masked_fill(mask[None], float("-inf"))
    p = torch.softmax(scores, dim=-1, dtype=torch.float32)
    return torch.einsum("hqk,khd->qhd", p, vv)


def swiglu(gate: torch.Tensor, up: torch.Tensor) -> torch.Tensor:
    """modeling_llama.py:182-184 — act_fn(gate) * up with SiLU."""
    return torch.nn.functional.silu(gate) * up


def bf16_round(x: torch.Tensor) -> torch.Tensor:
    return x.to(torch.bfloat16).to(torch.float32)


# --------------------------------------------------------------------------------------------------
# paged KV helpers (integer indexing must be bit-exact)
# --------------------------------------------------------------------------------------------------
def slot_mapping(block_table: List[int], start: int, count: int, block_size: int) -> List[int]:
    """slot = block_id * block_size + offset for token positions [start, start + count)."""
    return [block_table[p // block_size] * block_size + p % block_size for p in range(start, start + count)]


def gather_pages(cache: torch.Tensor, block_table: List[int], length: int) -> torch.Tensor:
    """cache: [num_blocks, Hkv, block_size, D] -> contiguous [length, Hkv, D]."""
    bs = cache.shape[2]
    rows = [cache[block_table[p // bs], :, p % bs, :] for p in range(length)]
    return torch.stack(rows, dim=0)


# --------------------------------------------------------------------------------------------------
# whole model
# --------------------------------------------------------------------------------------------------
class RefLlama:
    """Decoder-only forward with a contiguous KV cache (modeling_llama.py:303-333, 375-427, 445-499).

    `gemm_bf16_inputs=True` rounds every GEMM *input* activation to bf16 (weights already are): the
    arithmetic contract of the CUDA path (bf16 operands, fp32 accumulate, fp32 residual). With False
    it is the exact fp32 HF computation on the same weights.
    """

    def __init__(self, shape: LlamaShape, weights: Dict[str, torch.Tensor], gemm_bf16_inputs: bool = False,
                 kv_bf16: bool = False):
        self.s = shape
        self.w = weights
        self.inv_freq = llama3_inv_freq(shape)
        self.round_in = bf16_round if gemm_bf16_inputs else (lambda t: t)
        self.round_kv = bf16_round if kv_bf16 else (lambda t: t)
        self.k: List[Optional[torch.Tensor]] = [None] * shape.num_hidden_layers
        self.v: List[Optional[torch.Tensor]] = [None] * shape.num_hidden_layers
        self.pos = 0

    def reset(self) -> None:
        self.k = [None] * self.s.num_hidden_layers
        self.v = [None] * self.s.num_hidden_layers
        self.pos = 0

    def _lin(self, x: torch.Tensor, name: str) -> torch.Tensor:
        return self.round_in(x) @ self.w[name].T

    def forward(self, ids: torch.Tensor, return_hidden: bool = False) -> torch.Tensor:
        """ids: [T] int64 appended to the cache; returns logits [T, vocab] (fp32)."""
        s = self.s
        T = ids.shape[0]
        Hq, Hkv, D = s.num_attention_heads, s.num_key_value_heads, s.head_dim
        h = self.w["model.embed_tokens.weight"][ids]
        positions = torch.arange(self.pos, self.pos + T)
        cos, sin = rope_cos_sin(positions, self.inv_freq)
        for i in range(s.num_hidden_layers):
            p = f"model.layers.{i}."
            x = rms_norm(h, self.w[p + "input_layernorm.weight"], s.rms_norm_eps)
            q = self._lin(x, p + "self_attn.q_proj.weight").view(T, Hq, D)
            k = self._lin(x, p + "self_attn.k_proj.weight").view(T, Hkv, D)
            v = self._lin(x, p + "self_attn.v_proj.weight").view(T, Hkv, D)
            q = self.round_kv(apply_rope(q, cos, sin))
            k = self.round_kv(apply_rope(k, cos, sin))
            v = self.round_kv(v)
            self.k[i] = k if self.k[i] is None else torch.cat((self.k[i], k), 0)
            self.v[i] = v if self.v[i] is None else torch.cat((self.v[i], v), 0)
            a = attention(q, self.k[i], self.v[i], self.pos).reshape(T, Hq * D)
            h = h + self._lin(a, p + "self_attn.o_proj.weight")
            x = rms_norm(h, self.w[p + "post_attention_layernorm.weight"], s.rms_norm_eps)
            act = swiglu(self._lin(x, p + "mlp.gate_proj.weight"), self._lin(x, p + "mlp.up_proj.weight"))
            h = h + self._lin(act, p + "mlp.down_proj.weight")
        self.pos += T
        if return_hidden:
            return h
        x = rms_norm(h, self.w["model.norm.weight"], s.rms_norm_eps)
        head = "model.embed_tokens.weight" if s.tie_word_embeddings else "lm_head.weight"
        return self._lin(x, head)

    def greedy(self, prompt: torch.Tensor, new_tokens: int) -> Tuple[List[int], torch.Tensor]:
        """Greedy continuation (torch.argmax). Returns (ids, logits of every generated step)."""
        self.reset()
        logits = self.forward(prompt)[-1]
        out, steps = [], []
        for _ in range(new_tokens):
            steps.append(logits)
            nxt = int(torch.argmax(logits))
            out.append(nxt)
            logits = self.forward(torch.tensor([nxt]))[-1]
        return out, torch.stack(steps)
