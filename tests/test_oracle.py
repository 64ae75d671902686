"""The CPU oracle against the golden vectors minted from HF transformers (the reference's CPU
arithmetic, llm/hf_cpu_server.py:39-50) — see oracle/gen_golden.py."""
from pathlib import Path

import numpy as np
import pytest
import torch

from oracle import llama_ref as R
from oracle import sampler_ref as SR
from oracle.kv_ref import RefPagedKV, capacity_gauges

GOLD = Path(__file__).parent / "golden"


@pytest.mark.parametrize("name,seed", [("tiny", 11), ("tiny128", 21), ("tiny128_long", 21), ("tiny_long", 11)])
def test_restatement_reproduces_hf_golden(name, seed):
    g = np.load(GOLD / f"hf_{name}.npz")
    name = name.replace("_long", "")   # long prompts: 232 tokens on the head_dim-128 fixture (what smoke() runs), 200 on tiny
    shape = R.shape_by_name(name)
    w = R.init_weights(shape, seed)
    ref = R.RefLlama(shape, w)
    prompt = torch.from_numpy(g["prompt"]).long()
    with torch.no_grad():
        logits = ref.forward(prompt)
    assert np.abs(logits[-1].numpy() - g["logits_last"]).max() <= 1e-5
    assert np.abs(logits[0].numpy() - g["logits_first"]).max() <= 1e-5
    ids, _ = ref.greedy(prompt, 32)
    assert ids == g["greedy"].tolist()
    assert np.array_equal(R.llama3_inv_freq(shape).numpy(), g["inv_freq"])
    cos, sin = R.rope_cos_sin(torch.from_numpy(g["far_pos"]), R.llama3_inv_freq(shape))
    assert np.abs(cos.numpy() - g["far_cos"]).max() <= 1e-6 and np.abs(sin.numpy() - g["far_sin"]).max() <= 1e-6
    tab = R.cos_sin_table(shape, 8200)
    D = shape.head_dim
    assert torch.equal(tab[8192, : D // 2], cos[4, : D // 2]) and torch.equal(tab[8192, D // 2:], sin[4, : D // 2])


def test_llama3_rope_scaling_regimes():
    s = R.shape_by_name("8b")
    base = 1.0 / (s.rope_theta ** (torch.arange(0, 128, 2).float() / 128))
    f = R.llama3_inv_freq(s)
    wl = 2 * torch.pi / base
    assert torch.equal(f[wl < 8192 / 4], base[wl < 8192 / 4])            # high frequencies untouched
    assert torch.allclose(f[wl > 8192], base[wl > 8192] / 8)              # low frequencies / factor
    mid = (wl >= 8192 / 4) & (wl <= 8192)
    assert mid.any() and torch.all(f[mid] <= base[mid]) and torch.all(f[mid] >= base[mid] / 8)


def test_product_rope_table_equals_oracle_table():
    from atx.config import model_config_by_name
    from atx.weights import rope_table

    for name in ("1b", "8b", "tiny"):
        assert torch.equal(rope_table(model_config_by_name(name), 9000, torch.device("cpu")),
                           R.cos_sin_table(R.shape_by_name(name), 9000))


def test_shape_constants_match_survey():
    from atx.config import model_config_by_name

    m = model_config_by_name("8b")
    assert abs(m.weight_bytes_per_decode_step() / 1e9 - 15.01) < 0.01 and m.kv_bytes_per_token() == 131072
    assert abs(model_config_by_name("1b").weight_bytes_per_decode_step() / 1e9 - 2.47) < 0.01
    assert model_config_by_name("70b").kv_bytes_per_token() == 327680
    for n in ("1b", "8b", "70b", "tiny", "tiny128"):
        a, b = model_config_by_name(n), R.shape_by_name(n)
        for k in ("vocab_size", "hidden_size", "intermediate_size", "num_hidden_layers", "num_attention_heads",
                  "num_key_value_heads", "head_dim", "rope_theta", "tie_word_embeddings"):
            assert getattr(a, k) == getattr(b, k), (n, k)
        assert a.rope_scaling == b.rope_scaling


def test_attention_restatement_edge_cases():
    g = torch.Generator().manual_seed(0)
    q, k, v = (torch.randn(5, 4, 8, generator=g), torch.randn(9, 2, 8, generator=g), torch.randn(9, 2, 8, generator=g))
    out = R.attention(q, k, v, 4)
    # last query sees every key; first query sees keys 0..4 only
    full = torch.softmax(torch.einsum("hd,khd->hk", q[4], k.repeat_interleave(2, 1)) * 8 ** -0.5, -1)
    assert torch.allclose(out[4], torch.einsum("hk,khd->hd", full, v.repeat_interleave(2, 1)), atol=1e-6)
    part = R.attention(q[:1], k[:5], v[:5], 4)
    assert torch.allclose(out[0], part[0], atol=1e-6)


def test_sampler_restatement():
    logits = np.array([0.1, 3.0, 3.0, -1.0], dtype=np.float32)
    assert SR.sample(logits, 0.0, 1.0, 1, 0) == 1  # first maximum
    keep = SR.nucleus_mask(np.log(np.array([0.5, 0.3, 0.15, 0.05])), 1.0, 0.79)
    assert keep.tolist() == [True, True, False, False]
    keep = SR.nucleus_mask(np.log(np.array([0.5, 0.3, 0.15, 0.05])), 1.0, 0.81)
    assert keep.tolist() == [True, True, True, False]
    gum = SR.gumbel(123, 4, 100000)
    assert abs(gum.mean() - 0.5772) < 0.02 and abs(gum.var() - np.pi ** 2 / 6) < 0.05
    draws = [SR.sample(np.log(np.array([0.7, 0.2, 0.1], dtype=np.float32)), 1.0, 1.0, s, 0) for s in range(4000)]
    freq = np.bincount(draws, minlength=3) / 4000
    assert np.abs(freq - [0.7, 0.2, 0.1]).max() < 0.03


def test_kv_reference_model_and_capacity_formula():
    kv = RefPagedKV(4, 16)
    kv.grow("a", 17)
    kv.grow("b", 1)
    assert kv.tables == {"a": [0, 1], "b": [2]} and kv.slot("a", 16) == 16 and kv.slot("b", 0) == 32
    with pytest.raises(MemoryError):
        kv.grow("b", 40)
    kv.release("a")
    kv.grow("b", 33)
    assert kv.tables["b"] == [2, 3, 1]
    assert capacity_gauges(1000, 16, 4096) == (16000, 3, 16000 / 4096)
