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


def _tp_rank_main(rank, world, port, out_q):
    """One tensor-parallel rank on CPU: the engine's layer schedule (atx/runner.py:_forward) in oracle arithmetic
    on THIS rank's product-sharded weights, with the exchange step done the way the two-shot kernels do it —
    rank r owns the token rows t = r (mod world): reduce the partial o_proj / down_proj rows it owns, add its
    residual rows, normalise, all-gather the normalised rows — over torch.distributed `gloo`."""
    import os

    import torch.distributed as dist

    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        shape, cfg = R.shape_by_name("tiny"), model_config_by_name("tiny")
        sd = R.init_weights(shape, 11)
        w = ModelWeights.from_state_dict(cfg, sd, torch.device("cpu"), world, rank)
        T, H, D = 9, cfg.hidden_size, cfg.head_dim
        hq, hkv = cfg.num_attention_heads // world, cfg.num_key_value_heads // world
        ids = torch.randint(0, cfg.vocab_size, (T,), generator=torch.Generator().manual_seed(5))
        cos, sin = R.rope_cos_sin(torch.arange(T), R.llama3_inv_freq(shape))
        mine = torch.arange(T) % world == rank

        def exchange(h, part, ln):
            """all-reduce + residual add + RMSNorm, sharded by token row; h keeps only this rank's rows current."""
            rows = []
            for owner in range(world):      # reduce-scatter: the owner receives the sum of its rows
                sel = torch.arange(T) % world == owner
                buf = part[sel].clone()
                dist.reduce(buf, dst=owner, op=dist.ReduceOp.SUM)
                rows.append(buf)
            h[mine] += rows[rank]
            normed = R.rms_norm(h[mine], ln.float(), cfg.rms_norm_eps)
            n_max = (T + world - 1) // world
            padded = torch.zeros(n_max, H)
            padded[: normed.shape[0]] = normed
            gathered = [torch.zeros(n_max, H) for _ in range(world)]
            dist.all_gather(gathered, padded)
            xn = torch.zeros(T, H)
            for owner in range(world):
                sel = torch.arange(T) % world == owner
                xn[sel] = gathered[owner][: int(sel.sum())]
            return xn

        h = w.embed.float()[ids]
        xn = R.rms_norm(h, w.layers[0].ln1.float(), cfg.rms_norm_eps)       # first norm: local, all rows
        for li, l in enumerate(w.layers):
            qkv = xn @ l.wqkv.float().T
            q = R.apply_rope(qkv[:, : hq * D].view(T, hq, D), cos, sin)
            k = R.apply_rope(qkv[:, hq * D:(hq + hkv) * D].view(T, hkv, D), cos, sin)
            v = qkv[:, (hq + hkv) * D:].view(T, hkv, D)
            attn = R.attention(q, k, v, 0).reshape(T, hq * D)              # this rank's heads only
            xn = exchange(h, attn @ l.wo.float().T, l.ln2)
            v64 = l.wgu.float().reshape(-1, 2, 64, H)
            act = R.swiglu(xn @ v64[:, 0].reshape(-1, H).T, xn @ v64[:, 1].reshape(-1, H).T)
            nxt = w.layers[li + 1].ln1 if li + 1 < len(w.layers) else w.norm
            xn = exchange(h, act @ l.wdown.float().T, nxt)
        part_logits = xn @ w.lm_head.float().T                              # vocabulary shard
        all_logits = [torch.zeros_like(part_logits) for _ in range(world)]
        dist.all_gather(all_logits, part_logits)
        # by value (numpy): a torch tensor would travel as a shared-memory handle that dies with this process
        out_q.put((rank, torch.cat(all_logits, dim=1).numpy()))
    finally:
        dist.destroy_process_group()


def test_tp_data_flow_gloo_world2_matches_unsharded_oracle():
    import socket

    import torch.multiprocessing as tmp

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = tmp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_tp_rank_main, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = {r: torch.from_numpy(a) for r, a in (q.get(timeout=180) for _ in procs)}
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    shape = R.shape_by_name("tiny")
    ref = R.RefLlama(shape, R.init_weights(shape, 11))
    ids = torch.randint(0, shape.vocab_size, (9,), generator=torch.Generator().manual_seed(5))
    with torch.no_grad():
        want = ref.forward(ids)
    assert torch.equal(got[0], got[1])                       # every rank assembles the same logits
    assert float((got[0] - want).abs().max()) <= 2e-4        # fp32 everywhere: only the summation order differs
