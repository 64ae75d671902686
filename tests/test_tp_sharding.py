"""Tensor-parallel weight sharding on CPU: the shards of every rank recombine to the full-model result
(Megatron layout: q/k/v/gate/up by output rows, o/down by input columns, LM head by vocabulary)."""
import torch

from atx.config import model_config_by_name
from atx.weights import ModelWeights, interleave_gate_up
from oracle import llama_ref as R


def _deinterleave(wgu):
    two_i, h = wgu.shape
    v = wgu.reshape(two_i // 128, 2, 64, h)
    return v[:, 0].reshape(two_i // 2, h), v[:, 1].reshape(two_i // 2, h)


def test_shards_recombine_to_the_full_layer():
    shape, cfg = R.shape_by_name("tiny"), model_config_by_name("tiny")
    sd = R.init_weights(shape, 3)
    cpu = torch.device("cpu")
    full = ModelWeights.from_state_dict(cfg, sd, cpu)
    tp = 2
    shards = [ModelWeights.from_state_dict(cfg, sd, cpu, tp, r) for r in range(tp)]
    g = torch.Generator().manual_seed(0)
    T, H, D = 5, cfg.hidden_size, cfg.head_dim
    Hq, Hkv, I = cfg.num_attention_heads, cfg.num_key_value_heads, cfg.intermediate_size
    x = torch.randn(T, H, generator=g)
    lf = full.layers[0]
    qkv_full = x @ lf.wqkv.float().T
    q_full = qkv_full[:, : Hq * D].view(T, Hq, D)
    k_full = qkv_full[:, Hq * D:(Hq + Hkv) * D].view(T, Hkv, D)
    attn = torch.randn(T, Hq * D, generator=g)
    gate_f, up_f = _deinterleave(lf.wgu.float())
    act_full = R.swiglu(x @ gate_f.T, x @ up_f.T)
    o_sum = torch.zeros(T, H)
    d_sum = torch.zeros(T, H)
    logits = []
    for r, s in enumerate(shards):
        l = s.layers[0]
        qkv = x @ l.wqkv.float().T
        hq, hkv = Hq // tp, Hkv // tp
        # rank r owns query heads [r*hq, (r+1)*hq) and the kv heads they attend to
        assert torch.equal(qkv[:, : hq * D].view(T, hq, D), q_full[:, r * hq:(r + 1) * hq])
        assert torch.equal(qkv[:, hq * D:(hq + hkv) * D].view(T, hkv, D), k_full[:, r * hkv:(r + 1) * hkv])
        o_sum += attn[:, r * hq * D:(r + 1) * hq * D] @ l.wo.float().T
        gate, up = _deinterleave(l.wgu.float())
        act = R.swiglu(x @ gate.T, x @ up.T)
        assert torch.allclose(act, act_full[:, r * (I // tp):(r + 1) * (I // tp)], atol=1e-6)
        d_sum += act @ l.wdown.float().T
        logits.append(x @ s.lm_head.float().T)
        assert torch.equal(s.embed, full.embed) and torch.equal(s.norm, full.norm)
    assert torch.allclose(o_sum, attn @ lf.wo.float().T, atol=1e-4)
    assert torch.allclose(d_sum, act_full @ lf.wdown.float().T, atol=1e-4)
    assert torch.allclose(torch.cat(logits, dim=1), x @ full.lm_head.float().T, atol=1e-5)


def test_interleave_layout():
    gate = torch.arange(128 * 2, dtype=torch.float32).view(128, 2)
    up = -gate
    w = interleave_gate_up(gate, up)
    assert torch.equal(w[:64], gate[:64]) and torch.equal(w[64:128], up[:64])
    assert torch.equal(w[128:192], gate[64:]) and torch.equal(w[192:], up[64:])


def test_random_init_tp_shards_are_slices_of_the_single_gpu_model():
    """ModelWeights.random(tp_size=n) must give every rank the Megatron shard of the SAME model that
    random(tp_size=1) builds (bytes equal), replicated tensors identical on all ranks, so that tensor-parallel
    results can be checked against the unsharded engine and the ranks agree on embeddings and norms."""
    import torch

    from atx.config import model_config_by_name
    from atx.weights import ModelWeights

    for name in ("tiny", "tiny128"):
        cfg = model_config_by_name(name)
        dev = torch.device("cpu")
        full = ModelWeights.random(cfg, 7, dev)
        n = 2 if cfg.num_key_value_heads % 2 == 0 else 1
        if name == "tiny128":
            n = 1 if cfg.num_key_value_heads < 2 else 2
        if n == 1:
            continue
        D, Hq, Hkv, I = cfg.head_dim, cfg.num_attention_heads, cfg.num_key_value_heads, cfg.intermediate_size
        q_size, kv = Hq * D, Hkv * D
        for r in range(n):
            w = ModelWeights.random(cfg, 7, dev, tp_size=n, tp_rank=r)
            assert torch.equal(w.embed, full.embed) and torch.equal(w.norm, full.norm)
            for lf, ls in zip(full.layers, w.layers):
                assert torch.equal(ls.ln1, lf.ln1) and torch.equal(ls.ln2, lf.ln2)
                qs, ks = q_size // n, kv // n
                want_qkv = torch.cat([lf.wqkv[r * qs:(r + 1) * qs], lf.wqkv[q_size + r * ks: q_size + (r + 1) * ks],
                                      lf.wqkv[q_size + kv + r * ks: q_size + kv + (r + 1) * ks]])
                assert torch.equal(ls.wqkv, want_qkv)
                assert torch.equal(ls.wo, lf.wo[:, r * qs:(r + 1) * qs])
                rows = 2 * I // n
                assert torch.equal(ls.wgu, lf.wgu[r * rows:(r + 1) * rows])
                assert torch.equal(ls.wdown, lf.wdown[:, r * (I // n):(r + 1) * (I // n)])
            V = cfg.vocab_size // n
            assert torch.equal(w.lm_head, full.lm_head[r * V:(r + 1) * V])
