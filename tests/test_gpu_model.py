"""End-to-end parity of the engine (scheduler + paged KV + every kernel) against the CPU oracle on
identical bf16-rounded weights and prompts. Golden fixtures come from HF transformers
(tests/golden/hf_*.npz, minted by oracle/gen_golden.py)."""
from pathlib import Path

import numpy as np
import pytest
import torch

from oracle import llama_ref as R

pytestmark = pytest.mark.gpu
GOLD = Path(__file__).parent / "golden"


def _engine(shape_name, seed, **kw):
    from atx.config import EngineConfig, ModelSource, model_config_by_name
    from atx.engine import LLMEngine
    from atx.weights import ModelWeights

    shape = R.shape_by_name(shape_name)
    sd = R.init_weights(shape, seed)
    mcfg = model_config_by_name(shape_name)
    ecfg = EngineConfig(max_model_len=kw.pop("max_model_len", 256), max_num_seqs=kw.pop("max_num_seqs", 8),
                        max_num_batched_tokens=kw.pop("max_num_batched_tokens", 64),
                        num_gpu_blocks=kw.pop("num_gpu_blocks", 96),
                        async_scheduling=kw.pop("async_scheduling", False), **kw)   # tests read runner.logits per step
    w = ModelWeights.from_state_dict(mcfg, sd, torch.device("cuda", 0))
    return LLMEngine(ModelSource(mcfg, None, None), ecfg, w), shape, sd


def _run_collect(eng, prompt, n_new, temperature=0.0):
    """Generate n_new tokens for one request, returning (ids, logits row of every sampling step)."""
    from atx.scheduler import SamplingParams

    seq = eng.add_request("r0", list(prompt), SamplingParams(temperature=temperature, max_tokens=n_new,
                                                             ignore_eos=True))
    rows = []
    while eng.has_work():
        got = eng.step()
        if got:
            rows.append(eng.runner.logits[0].float().cpu().clone())
    return seq.output_ids, torch.stack(rows)


@pytest.mark.parametrize("shape_name,gold,seed", [("tiny", "hf_tiny.npz", 11), ("tiny128", "hf_tiny128.npz", 21),
                                                  ("tiny128", "hf_tiny128_long.npz", 21),
                                                  ("tiny", "hf_tiny_long.npz", 11)])
@pytest.mark.parametrize("graphs", [False, True])
def test_greedy_and_logits_against_hf_golden(shape_name, gold, seed, graphs):
    g = np.load(GOLD / gold)
    prompt = g["prompt"].tolist()
    # the 232-token prompt goes through in ONE chunk: prefill-sized kernels (persistent GEMM, two query tiles)
    big = dict(max_model_len=512, max_num_batched_tokens=256) if len(prompt) > 64 else {}
    eng, shape, sd = _engine(shape_name, seed, enable_cuda_graphs=graphs, **big)
    ids, logits = _run_collect(eng, prompt, 32)
    assert len(ids) == 32

    # (1) first-step logits vs the HF fp32 golden vector (literal north_star criterion: 2e-2 max-abs)
    err_first = float((logits[0] - torch.from_numpy(g["logits_last"])).abs().max())
    # (2) teacher-forced: oracle replays OUR tokens; compare every step's logits
    exact = R.RefLlama(shape, sd)
    emul = R.RefLlama(shape, sd, gemm_bf16_inputs=True, kv_bf16=True)
    seq_ids = torch.tensor(prompt + ids[:-1])
    P = len(prompt)
    with torch.no_grad():
        lo_exact = exact.forward(seq_ids)[P - 1:]
        lo_emul = emul.forward(seq_ids)[P - 1:]
    err_exact = (logits - lo_exact).abs().max(dim=-1).values
    err_emul = (logits - lo_emul).abs().max(dim=-1).values
    hf_greedy = g["greedy"].tolist()
    n_same = next((i for i, (a, b) in enumerate(zip(ids, hf_greedy)) if a != b), 32)
    print(f"\n[{shape_name} graphs={graphs}] first-step max-abs vs HF fp32 = {err_first:.4f}; "
          f"teacher-forced max-abs vs fp32 oracle = {float(err_exact.max()):.4f}, vs bf16-operand oracle = "
          f"{float(err_emul.max()):.5f}; greedy identical to HF for {n_same}/32 tokens")
    # against the oracle that rounds the same operands to bf16 only fp32 summation order differs
    # (a bf16 rounding decision flips wherever the fp32 sums differ in their last bits: 2^-8 relative)
    assert float(err_emul.max()) <= 1.5e-2
    # literal criterion of the north star on this shape
    assert err_first <= 2e-2 and float(err_exact.max()) <= 2e-2
    # arg-max agreement wherever the fp32 oracle's top-1/top-2 margin exceeds 4x the measured error
    top2 = lo_exact.topk(2, dim=-1).values
    margin = top2[:, 0] - top2[:, 1]
    safe = margin > 4 * err_exact.max()
    assert int(safe.sum()) >= 16, "fixture should have enough decisive positions"
    assert torch.equal(logits.argmax(-1)[safe], lo_exact.argmax(-1)[safe])
    if bool(safe.all()):
        assert ids == hf_greedy


def test_kv_pages_match_oracle_cache():
    """Pages read back through the block table equal the oracle's contiguous K/V (post-RoPE, bf16):
    indexing bit-exact, values within one bf16 ulp (fp32 summation order is the only difference)."""
    g = np.load(GOLD / "hf_tiny.npz")
    prompt = g["prompt"].tolist()
    eng, shape, sd = _engine("tiny", 11)
    from atx.scheduler import SamplingParams

    seq = eng.add_request("kv", prompt, SamplingParams(temperature=0.0, max_tokens=4, ignore_eos=True))
    eng.step()  # 40-token prompt, budget 64: one chunk
    table = list(seq.blocks.table)
    P = len(prompt)
    assert seq.blocks.slots(0, P) == R.slot_mapping(table, 0, P, 16)
    emul = R.RefLlama(shape, sd, gemm_bf16_inputs=True, kv_bf16=True)
    with torch.no_grad():
        emul.forward(torch.tensor(prompt))
    for li in range(shape.num_hidden_layers):
        k = R.gather_pages(eng.runner.k_cache[li].float().cpu(), table, P)
        v = R.gather_pages(eng.runner.v_cache[li].float().cpu(), table, P)
        for name, got, want in (("k", k, emul.k[li]), ("v", v, emul.v[li])):
            d = (got - want).abs()
            print(f"layer {li} {name}: max abs {float(d.max()):.3e}, max rel-to-ulp "
                  f"{float((d / (2 ** -7 * want.abs() + 1e-6)).max()):.2f}, bit-equal {float((got == want).float().mean()):.4f}")
            if li == 0:  # same inputs on both sides: only the fp32 summation order differs
                assert torch.all((got - want).abs() <= 2 ** -7 * want.abs() + 1e-6)
                assert float((got == want).float().mean()) > 0.98
            else:        # deeper layers inherit flipped bf16 roundings of the layers below
                assert torch.all((got - want).abs() <= 2 ** -5 * want.abs() + 2e-2)


def _decisive_agreement(shape, sd, prompt, ids, tol=0.05):
    """Teacher-forced check of generated ids against the fp32 oracle: wherever the oracle's top-1/top-2
    margin exceeds `tol` (several times the measured logit error) the id must be the oracle's arg-max.
    Returns (#decisive positions, #positions). Batch composition changes fp32 summation order (tile
    size, split-K), so bit-identical tokens across schedules are not a valid expectation on a
    random-init model whose margins go down to 0.03."""
    ref = R.RefLlama(shape, sd)
    with torch.no_grad():
        lo = ref.forward(torch.tensor(list(prompt) + list(ids[:-1])))[len(prompt) - 1:]
    top2 = lo.topk(2, dim=-1).values
    decisive = (top2[:, 0] - top2[:, 1]) > tol
    got = torch.tensor(ids)
    assert torch.equal(got[decisive], lo.argmax(-1)[decisive]), (got.tolist(), lo.argmax(-1).tolist())
    return int(decisive.sum()), len(ids)


def test_batched_chunked_prefix_cached_equals_single():
    """6 requests sharing a 96-token prefix (the agent fan-out shape), admitted together under a
    64-token budget (chunked prefill + continuous batching + in-burst prefix sharing), and each
    request run alone without the prefix cache: both must follow the oracle at every decisive
    position, and the prefix cache must actually have been used."""
    from atx.scheduler import SamplingParams

    g = torch.Generator().manual_seed(5)
    prefix = torch.randint(0, 1024, (96,), generator=g).tolist()
    prompts = [prefix + torch.randint(0, 1024, (7 + 5 * i,), generator=g).tolist() for i in range(6)]
    params = SamplingParams(temperature=0.0, max_tokens=12, ignore_eos=True)

    eng, shape, sd = _engine("tiny", 11, num_gpu_blocks=128)
    first = eng.add_request("warm", prompts[0], params)
    while eng.has_work():
        eng.step()
    seqs = [eng.add_request(f"b{i}", p, params) for i, p in enumerate(prompts)]
    while eng.has_work():
        eng.step()
    assert all(s.cached_prompt_tokens >= 96 for s in seqs), [s.cached_prompt_tokens for s in seqs]
    assert eng.allocator.num_used == 0  # everything released
    batched = [s.output_ids for s in seqs]
    n_dec = n_all = same = 0
    for i, p in enumerate(prompts):
        d, n = _decisive_agreement(shape, sd, p, batched[i])
        n_dec, n_all = n_dec + d, n_all + n
        solo, _, _ = _engine("tiny", 11, enable_prefix_caching=False, max_num_batched_tokens=256)
        out = solo.generate_sync([p], params)[0]
        _decisive_agreement(shape, sd, p, out)
        same += out == batched[i]
    _decisive_agreement(shape, sd, prompts[0], first.output_ids)
    print(f"\ndecisive positions {n_dec}/{n_all}; batched == solo for {same}/6 requests")
    assert n_dec >= n_all // 2 and same >= 3


def test_preemption_recompute_is_transparent():
    """With a KV pool too small for all sequences the scheduler preempts and recomputes; outputs
    must equal the unconstrained run."""
    from atx.scheduler import SamplingParams

    g = torch.Generator().manual_seed(6)
    prompts = [torch.randint(0, 1024, (40,), generator=g).tolist() for _ in range(4)]
    params = SamplingParams(temperature=0.0, max_tokens=40, ignore_eos=True)
    big, _, _ = _engine("tiny", 11, num_gpu_blocks=128, enable_prefix_caching=False)
    want = big.generate_sync(prompts, params)
    small, _, _ = _engine("tiny", 11, num_gpu_blocks=16, enable_prefix_caching=False)  # 256 tokens for 4 x 80
    seqs = [small.add_request(f"p{i}", p, params) for i, p in enumerate(prompts)]
    while small.has_work():
        small.step()
    assert sum(s.num_preemptions for s in seqs) > 0
    shape, sd = R.shape_by_name("tiny"), R.init_weights(R.shape_by_name("tiny"), 11)
    for p, s, w in zip(prompts, seqs, want):
        _decisive_agreement(shape, sd, p, s.output_ids)
        _decisive_agreement(shape, sd, p, w)
    assert sum(s.output_ids == w for s, w in zip(seqs, want)) >= 2


def test_async_scheduling_produces_the_same_tokens():
    """Launching step N+1 before the host has read step N's tokens (device-side token feed) must not
    change anything: same greedy ids as the synchronous engine, EOS / length stops included."""
    from atx.scheduler import SamplingParams

    g = torch.Generator().manual_seed(9)
    prompts = [torch.randint(0, 1000, (20 + 7 * i,), generator=g).tolist() for i in range(5)]
    lens = [3, 17, 40, 41, 64]
    outs = []
    for mode in (False, True):
        eng, _, _ = _engine("tiny", 11, num_gpu_blocks=128, async_scheduling=mode, max_num_batched_tokens=256)
        eng.scheduler.eos = {7}   # make EOS stops happen on this random model
        seqs = [eng.add_request(f"a{i}", p, SamplingParams(temperature=0.0, max_tokens=n)) for i, (p, n) in
                enumerate(zip(prompts, lens))]
        late = None
        steps = 0
        while eng.has_work():
            eng.step()
            steps += 1
            if steps == 10:   # a request that arrives while a chained step is in flight
                late = eng.add_request("late", prompts[0][:9], SamplingParams(temperature=0.0, max_tokens=5))
        outs.append([s.output_ids for s in seqs] + [late.output_ids])
        assert eng.allocator.num_used == 0 and eng._inflight is None
        assert all(s.finish_reason in ("stop", "length") for s in seqs)
    assert outs[0] == outs[1]


def test_deterministic_mode_repeats_itself_bit_for_bit():
    """EngineConfig.deterministic (LLM_DETERMINISTIC=1): the same burst, run twice on fresh engines through chunked
    prefill, prefix-cache hits and graph-replayed decode steps, yields bit-identical logits and identical tokens.
    (Without it the split-K adds race, and a last-bit difference may flip a near-tie: see DESIGN.md section 5.)"""
    g = torch.Generator().manual_seed(9)
    shared = torch.randint(0, 1000, (48,), generator=g).tolist()
    prompts = [shared + torch.randint(0, 1000, (10 + 7 * i,), generator=g).tolist() for i in range(6)]
    from atx.scheduler import SamplingParams

    outs = []
    for rep in range(2):
        eng, _, _ = _engine("tiny128", 21, deterministic=True, enable_cuda_graphs=True, max_num_batched_tokens=128)
        seqs = [eng.add_request(f"r{i}", p, SamplingParams(temperature=0.0 if i % 2 else 0.7, seed=5 + i, max_tokens=24,
                                                           ignore_eos=True)) for i, p in enumerate(prompts)]
        rows = []
        while eng.has_work():
            if eng.step():
                rows.append(eng.runner.logits[:6].clone())
        outs.append(([s.output_ids for s in seqs], torch.stack(rows).cpu()))
    assert outs[0][0] == outs[1][0]
    assert torch.equal(outs[0][1].view(torch.int32), outs[1][1].view(torch.int32))


def test_no_cuda_means_loud_failure(monkeypatch):
    from atx.config import EngineConfig, model_config_by_name
    from atx.runner import ModelRunner

    monkeypatch.setattr(torch.cuda, "is_available", lambda: False)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ModelRunner(model_config_by_name("tiny"), EngineConfig(), None)


_FULL_DEPTH = pytest.mark.skipif(__import__("os").environ.get("ATX_RUN_SLOW") != "1",
                                 reason="full-depth 8B: 32 GB of fp32 oracle weights and minutes of CPU time; "
                                        "run once on the box with ATX_RUN_SLOW=1 (numbers recorded in DESIGN.md)")


@pytest.mark.parametrize("shape_name,layers", [("1b", 2), ("8b", 1), ("70b", 1),
                                               pytest.param("1b", 16, marks=[pytest.mark.slow, _FULL_DEPTH]),
                                               pytest.param("8b", 32, marks=[pytest.mark.slow, _FULL_DEPTH])])
def test_baseline_shapes_full_width_reduced_depth(shape_name, layers):
    """The BASELINE.json model shapes at FULL width — hidden size, GQA head counts (group 4; group 8 with 64 query
    heads, hidden 8192 and a 28 672-wide MLP for the 70B shape), head_dim 64 (1B) / 128 (8B, 70B),
    the 128 256-token vocabulary (1002 LM-head tiles, the sampler over a real row), tied (1B) / untied (8B) head,
    llama3 rope scaling x32 / x8 — at reduced depth so the fp32 oracle finishes in seconds, on
    BASELINE configs[0]'s input: batch 1, a 128-token prompt of ids below 128 000, greedy decoding.
    Gates: logits within bf16-rounding distance of the oracle that rounds the same operands; arg-max equal to
    the fp32 oracle's wherever its top-1/top-2 margin is decisive; slot mapping bit-exact."""
    from atx.config import EngineConfig, ModelSource, model_config_by_name
    from atx.engine import LLMEngine
    from atx.weights import ModelWeights

    shape = R.shape_by_name(shape_name, layers)
    mcfg = model_config_by_name(shape_name, layers)
    sd = R.init_weights(shape, 5)
    w = ModelWeights.from_state_dict(mcfg, sd, torch.device("cuda", 0))
    eng = LLMEngine(ModelSource(mcfg, None, None),
                    EngineConfig(max_model_len=512, max_num_seqs=4, max_num_batched_tokens=256, num_gpu_blocks=64,
                                 async_scheduling=False), w)
    g = torch.Generator().manual_seed(1)
    prompt = torch.randint(0, 128000, (128,), generator=g).tolist()
    full_depth = layers >= 16
    n_new = 32 if full_depth else 8   # full depth (SURVEY 8c "once on the box"): the north star's 32 tokens
    ids, logits = _run_collect(eng, prompt, n_new)
    assert logits.shape == (n_new, 128256)
    exact = R.RefLlama(shape, sd)
    emul = R.RefLlama(shape, sd, gemm_bf16_inputs=True, kv_bf16=True)
    seq_ids = torch.tensor(prompt + ids[:-1])
    P = len(prompt)
    with torch.no_grad():
        lo_exact = exact.forward(seq_ids)[P - 1:]
        lo_emul = emul.forward(seq_ids)[P - 1:]
    err_exact = float((logits - lo_exact).abs().max())
    err_emul = float((logits - lo_emul).abs().max())
    hf_bf16_vs_fp32 = float((lo_emul - lo_exact).abs().max())   # what ANY bf16-operand pipeline loses on these weights
    scale = float(lo_exact.abs().max())
    top2 = lo_exact.topk(2, dim=-1).values
    margin = top2[:, 0] - top2[:, 1]
    safe = margin > 4 * err_exact
    print(f"\n[{shape_name} x{layers} layers, full width] logits scale {scale:.2f}; max-abs vs bf16-operand oracle "
          f"{err_emul:.4f}, vs fp32 oracle {err_exact:.4f} (the bf16-operand oracle itself: {hf_bf16_vs_fp32:.4f}); "
          f"decisive positions {int(safe.sum())}/{n_new}, min margin {float(margin.min()):.3f}")
    # Over 8 x 128 256 logits of a 2048/4096-wide model the yardstick is the noise of the arithmetic contract
    # itself: what rounding GEMM operands to bf16 costs against fp32 on these very weights (0.035 on the 1B
    # shape). Our deviation from the bf16-operand oracle is of the same nature (fp32 summation order flips bf16
    # roundings of intermediate activations) and must not exceed it by much; nor may the total.
    assert err_emul <= 1.5 * hf_bf16_vs_fp32 + 1e-2, (err_emul, hf_bf16_vs_fp32)
    assert err_exact <= 2.0 * hf_bf16_vs_fp32 + 1e-2, (err_exact, hf_bf16_vs_fp32)
    if full_depth:
        with torch.no_grad():
            hf_greedy, _ = R.RefLlama(shape, sd).greedy(torch.tensor(prompt), n_new)
        n_same = next((i for i, (a, b) in enumerate(zip(ids, hf_greedy)) if a != b), n_new)
        rms = float((logits - lo_exact).pow(2).mean().sqrt())
        print(f"[{shape_name} x{layers}] literal north-star criteria: logits max-abs vs fp32 {err_exact:.4f} "
              f"(bound 2e-2: {'met' if err_exact <= 2e-2 else 'NOT met'}; rms {rms:.4f}); free-running greedy "
              f"identical to the fp32 oracle for {n_same}/{n_new} tokens; teacher-forced arg-max agreement "
              f"{int((logits.argmax(-1) == lo_exact.argmax(-1)).sum())}/{n_new}")
    assert torch.equal(logits.argmax(-1)[safe], lo_exact.argmax(-1)[safe])
    assert ids == logits.argmax(-1).tolist()                # the sampler's arg-max over the full 128 256-wide row
