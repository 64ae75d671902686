"""Loading a real HF-format checkpoint directory (what `--model /models/Llama-3.1-8B-Instruct` points at): the
directory is written by transformers' own `save_pretrained` (config.json + model.safetensors; tied heads are
NOT stored, as in the public 1B / 3B checkpoints), read back by atx.config.resolve_model +
ModelWeights.from_hf_dir, and must give the layout the kernels stream: fused q/k/v, gate/up interleaved in
64-row groups, bf16 bytes equal to the checkpoint's, Megatron shards under tensor parallelism. CPU only."""
import json

import pytest
import torch

from atx.config import resolve_model
from atx.weights import ModelWeights, interleave_gate_up


def _write_checkpoint(tmp_path, tie: bool):
    from transformers import LlamaConfig, LlamaForCausalLM

    cfg = LlamaConfig(vocab_size=512, hidden_size=256, intermediate_size=512, num_hidden_layers=2,
                      num_attention_heads=4, num_key_value_heads=2, head_dim=64, max_position_embeddings=4096,
                      rms_norm_eps=1e-5, rope_theta=500000.0, tie_word_embeddings=tie,
                      rope_scaling={"rope_type": "llama3", "factor": 8.0, "low_freq_factor": 1.0,
                                    "high_freq_factor": 4.0, "original_max_position_embeddings": 8192},
                      bos_token_id=1, eos_token_id=[2, 3], torch_dtype="bfloat16")
    torch.manual_seed(4)
    model = LlamaForCausalLM(cfg).to(torch.bfloat16)
    d = tmp_path / ("tied" if tie else "untied")
    model.save_pretrained(d, safe_serialization=True)
    return d, model


@pytest.mark.parametrize("tie", [False, True])
def test_hf_checkpoint_directory_round_trip(tmp_path, tie):
    d, model = _write_checkpoint(tmp_path, tie)
    assert list(d.glob("*.safetensors")) and (d / "config.json").exists()
    src = resolve_model(str(d))
    m = src.config
    raw = json.loads((d / "config.json").read_text())
    assert src.path == d and src.random_seed is None
    assert (m.vocab_size, m.hidden_size, m.intermediate_size, m.num_hidden_layers) == (512, 256, 512, 2)
    assert (m.num_attention_heads, m.num_key_value_heads, m.head_dim) == (4, 2, 64)
    assert m.tie_word_embeddings == tie and m.rms_norm_eps == raw["rms_norm_eps"]
    assert set(m.eos_token_ids) == {2, 3}
    sd = model.state_dict()
    dev = torch.device("cpu")
    w = ModelWeights.from_hf_dir(m, d, dev)
    assert torch.equal(w.embed, sd["model.embed_tokens.weight"])
    head = sd["model.embed_tokens.weight"] if tie else sd["lm_head.weight"]
    assert torch.equal(w.lm_head, head) and (w.lm_head is w.embed) == tie
    for i, l in enumerate(w.layers):
        p = f"model.layers.{i}."
        assert torch.equal(l.wqkv, torch.cat([sd[p + "self_attn.q_proj.weight"], sd[p + "self_attn.k_proj.weight"],
                                              sd[p + "self_attn.v_proj.weight"]]))
        assert torch.equal(l.wo, sd[p + "self_attn.o_proj.weight"])
        assert torch.equal(l.wgu, interleave_gate_up(sd[p + "mlp.gate_proj.weight"], sd[p + "mlp.up_proj.weight"]))
        assert torch.equal(l.wdown, sd[p + "mlp.down_proj.weight"])
        assert torch.equal(l.ln1, sd[p + "input_layernorm.weight"])
        assert torch.equal(l.ln2, sd[p + "post_attention_layernorm.weight"])
    assert all(t.dtype == torch.bfloat16 for t in w._all())
    # tensor parallel: every rank keeps its Megatron shard of the same checkpoint
    n = 2
    for r in range(n):
        s = ModelWeights.from_hf_dir(m, d, dev, tp_size=n, tp_rank=r)
        V = m.vocab_size // n
        assert torch.equal(s.lm_head, head[r * V:(r + 1) * V]) and torch.equal(s.embed, w.embed)
        for lf, ls in zip(w.layers, s.layers):
            qs, ks = m.q_size // n, m.kv_size // n
            assert torch.equal(ls.wqkv, torch.cat([lf.wqkv[r * qs:(r + 1) * qs],
                                                   lf.wqkv[m.q_size + r * ks: m.q_size + (r + 1) * ks],
                                                   lf.wqkv[m.q_size + m.kv_size + r * ks: m.q_size + m.kv_size + (r + 1) * ks]]))
            assert torch.equal(ls.wo, lf.wo[:, r * qs:(r + 1) * qs])
            rows = 2 * m.intermediate_size // n
            assert torch.equal(ls.wgu, lf.wgu[r * rows:(r + 1) * rows])
            assert torch.equal(ls.wdown, lf.wdown[:, r * (m.intermediate_size // n):(r + 1) * (m.intermediate_size // n)])


def test_sharded_checkpoint_is_read_slice_by_slice_and_decoys_are_ignored(tmp_path):
    """The hub layout of the large models: several model-0000x-of-0000N.safetensors shards + model.safetensors.index.json,
    often with Meta's `consolidated.*.safetensors` next to them. A tensor-parallel rank reads only the slices it
    keeps (host memory: one tensor at a time; it used to read every shard whole on every rank) and never opens a
    file the index does not name."""
    from safetensors.torch import save_file

    from atx.weights import _SafetensorsSource

    d, model = _write_checkpoint(tmp_path, False)
    sd = {k: v.contiguous() for k, v in model.state_dict().items()}
    (d / "model.safetensors").unlink()
    names = sorted(sd)
    parts = [names[0::3], names[1::3], names[2::3]]
    wmap = {}
    for i, part in enumerate(parts):
        fname = f"model-{i + 1:05d}-of-00003.safetensors"
        save_file({k: sd[k] for k in part}, str(d / fname))
        wmap.update({k: fname for k in part})
    (d / "model.safetensors.index.json").write_text(json.dumps({"metadata": {}, "weight_map": wmap}))
    (d / "consolidated.00.safetensors").write_bytes(b"not a safetensors file at all")     # must never be opened
    m = resolve_model(str(d)).config
    dev = torch.device("cpu")
    full = ModelWeights.from_hf_dir(m, d, dev)
    assert torch.equal(full.layers[1].wdown, sd["model.layers.1.mlp.down_proj.weight"])
    total = sum(v.numel() * 2 for v in sd.values())
    replicated = sum(v.numel() * 2 for k, v in sd.items() if "norm" in k or "embed" in k)
    for r in range(2):
        src = _SafetensorsSource(d)
        s = ModelWeights._from_source(m, src, dev, 2, r)
        assert src.bytes_read == replicated + (total - replicated) // 2     # its shard and nothing else
        assert torch.equal(s.layers[0].wo, full.layers[0].wo[:, r * (m.q_size // 2):(r + 1) * (m.q_size // 2)])
    # without an index the model*.safetensors files are scanned; the decoy still is not touched
    (d / "model.safetensors.index.json").unlink()
    again = ModelWeights.from_hf_dir(m, d, dev)
    assert torch.equal(again.layers[1].wgu, full.layers[1].wgu)


def test_missing_checkpoint_is_a_loud_error(tmp_path, monkeypatch):
    monkeypatch.delenv("LLM_RANDOM_INIT_SEED", raising=False)
    monkeypatch.setenv("HF_HUB_OFFLINE", "1")
    with pytest.raises(FileNotFoundError, match="neither a local directory nor in the HF cache"):
        resolve_model("meta-llama/Llama-3.1-8B-Instruct-not-here")
    (tmp_path / "config.json").write_text(json.dumps({"model_type": "llama"}))
    with pytest.raises(Exception):
        ModelWeights.from_hf_dir(resolve_model("random:tiny").config, tmp_path, torch.device("cpu"))


def _write_tokenizer(d):
    """A small byte-level BPE tokenizer in the HF on-disk format with the Llama-3 special tokens and a chat
    template in tokenizer_config.json (the real files cannot be fetched offline; the format is the same)."""
    from tokenizers import Tokenizer, decoders, models, pre_tokenizers, trainers
    from transformers import PreTrainedTokenizerFast

    tok = Tokenizer(models.BPE())
    tok.pre_tokenizer = pre_tokenizers.ByteLevel(add_prefix_space=False)
    tok.decoder = decoders.ByteLevel()
    specials = ["<|begin_of_text|>", "<|end_of_text|>", "<|start_header_id|>", "<|end_header_id|>", "<|eot_id|>"]
    trainer = trainers.BpeTrainer(vocab_size=400, special_tokens=specials,
                                  initial_alphabet=pre_tokenizers.ByteLevel.alphabet())
    tok.train_from_iterator(["the agent asks the tool for the weather in paris", "plan then act then review",
                             "hello world, hello agents"] * 20, trainer)
    fast = PreTrainedTokenizerFast(tokenizer_object=tok, bos_token="<|begin_of_text|>", eos_token="<|eot_id|>")
    fast.chat_template = ("{{ bos_token }}{% for m in messages %}<|start_header_id|>{{ m['role'] }}<|end_header_id|>\n\n"
                          "{{ m['content'] }}<|eot_id|>{% endfor %}{% if add_generation_prompt %}"
                          "<|start_header_id|>assistant<|end_header_id|>\n\nCKPT:{% endif %}")
    fast.save_pretrained(d)
    return fast


def test_checkpoint_tokenizer_and_its_chat_template_are_used(tmp_path):
    from atx.tokenizer import ByteTokenizer, HFTokenizer, load_tokenizer
    from llm.backend import B200Backend
    from tests.fakes import FakeAsyncEngine

    assert isinstance(load_tokenizer(tmp_path, 128256), ByteTokenizer)       # no tokenizer files: offline stub
    _write_tokenizer(tmp_path)
    tk = load_tokenizer(tmp_path, 128256)
    assert isinstance(tk, HFTokenizer)
    text = "the agent asks the tool"
    ids = tk.encode(text)
    assert ids and tk.decode(ids) == text and len(ids) < len(text)           # merges learnt: not byte-per-token
    eot = tk.tok.convert_tokens_to_ids("<|eot_id|>")
    assert tk.decode(ids + [eot], skip_special_tokens=True) == text          # `output` carries no special tokens
    be = B200Backend("ckpt", 256, None, 4, 64, 0.5, engine=FakeAsyncEngine(), tokenizer=tk)
    assert be.count_tokens(text) == len(ids)
    templ = be.apply_chat_template("hi", "be brief")
    # the checkpoint's own template wins over the literal fallback of llm/serve_llm.py:671-678
    assert templ.endswith("<|start_header_id|>assistant<|end_header_id|>\n\nCKPT:")
    assert "<|start_header_id|>system<|end_header_id|>\n\nbe brief<|eot_id|>" in templ
    assert "<|start_header_id|>user<|end_header_id|>\n\nhi<|eot_id|>" in templ
