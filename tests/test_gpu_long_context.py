"""Attention parity at the sizes BASELINE.json configs[4] runs (16 x 32 768-token contexts, 8 192-token prefill
chunks over a cached prefix, 16-32 KV splits, a 73 785-block pool) with a reduced head count so that the CPU
oracle (oracle/llama_ref.py:attention = modeling_llama.py:199-222) stays cheap. What these reach that the
short-context tests do not: thousands of 16-token tiles per split chain, the lazy-rescale branch of the tcgen05
prefill after hundreds of tiles, ex2 accumulation over 32k keys, TMA row coordinates near num_blocks * 16.
"""
import pytest
import torch

from oracle import llama_ref as R

pytestmark = pytest.mark.gpu

POOL = 73785          # blocks of the configs[4] pool (tests/test_kv_scheduler.py uses the same figure)
BS = 16


@pytest.fixture(scope="module")
def cabi():
    from atx import _cabi

    _cabi.load()
    return _cabi


def _scatter(kc, vc, tbl, k, v):
    """Write one sequence's K/V [L, Hkv, D] into the device pools through its block table (vectorised)."""
    L = k.shape[0]
    pos = torch.arange(L)
    blocks = torch.tensor(tbl, dtype=torch.long)[pos // BS].cuda()
    offs = (pos % BS).cuda()
    kc[blocks, :, offs] = k.cuda()
    vc[blocks, :, offs] = v.cuda()


def _pool(Hkv, D):
    """Pools of the full configs[4] block count, NaN everywhere a test did not write."""
    kc = torch.full((POOL, Hkv, BS, D), float("nan"), dtype=torch.bfloat16, device="cuda")
    vc = torch.full_like(kc, float("nan"))
    return kc, vc


def _tables(lens, gen, top=True):
    """Shuffled block ids taken from the TOP of the pool (the last ids the allocator hands out)."""
    need = [(L + BS - 1) // BS for L in lens]
    total = sum(need)
    ids = (torch.randperm(total, generator=gen) + (POOL - total if top else 0)).tolist()
    tables, o = [], 0
    for n in need:
        tables.append(ids[o:o + n])
        o += n
    width = max(need) + 1
    bt = torch.tensor([t + [0] * (width - len(t)) for t in tables], dtype=torch.int32)
    return tables, bt


@pytest.mark.parametrize("Hq,Hkv", [(4, 1), (8, 1)])      # GQA group 4 (8B) and 8 (70B), one KV head
@pytest.mark.parametrize("splits", [16, 32])
def test_decode_attention_at_32k_context_batch_16(cabi, Hq, Hkv, splits):
    D = 128
    g = torch.Generator().manual_seed(5 + Hq + splits)
    lens = [8191, 8192, 32768, 32896, 32767, 32769, 16384, 1, 32896, 8193, 24576, 32768, 17, 32895, 30000, 32896]
    B = len(lens)
    tables, bt = _tables(lens, g)
    kc, vc = _pool(Hkv, D)
    ks, vs = [], []
    for L, tbl in zip(lens, tables):
        k = torch.randn(L, Hkv, D, generator=g).to(torch.bfloat16)
        v = torch.randn(L, Hkv, D, generator=g).to(torch.bfloat16)
        _scatter(kc, vc, tbl, k, v)
        ks.append(k.float())
        vs.append(v.float())
    # scores of std ~3 (a few keys dominate, as in trained models) rather than ~1
    q = (3 * torch.randn(B, Hq, D, generator=g)).to(torch.bfloat16)
    out = torch.full((B, Hq * D), float("nan"), dtype=torch.bfloat16, device="cuda")
    ws = torch.empty(max(1, cabi.decode_workspace_bytes(B, Hq, D, splits) // 4), device="cuda")
    cabi.paged_attn_decode(q.cuda().view(B, Hq * D), kc, vc, bt.cuda(), torch.tensor(lens, dtype=torch.int32).cuda(),
                           out, ws, B, Hq, Hkv, D, BS, splits, D ** -0.5)
    got = out.float().cpu().view(B, Hq, D)
    assert torch.isfinite(got).all()
    worst = 0.0
    for b, L in enumerate(lens):
        want = R.attention(q[b].float()[None], ks[b], vs[b], L - 1)[0]
        err = (got[b] - want).abs()
        worst = max(worst, float(err.max()))
        # fp32 softmax / accumulation, one bf16 rounding of the output (same bound as the short-context test)
        assert torch.all(err <= 2 ** -7 * want.abs() + 2e-3), (b, L, float(err.max()))
    print(f"\n[decode 16 x <=32896, Hq={Hq}, splits={splits}] max-abs err {worst:.2e}")


def _check_windows(got, q, k, v, off, windows, tag):
    """Oracle on 128-row windows of the chunk (each query row is independent of the others)."""
    worst = 0.0
    for m in windows:
        hi = off + m + 128
        want = R.attention(q[m:m + 128].float(), k[:hi], v[:hi], off + m)
        err = (got[m:m + 128] - want).abs()
        worst = max(worst, float(err.max()))
        # P is rounded to bf16 before P.V (2^-9 relative per term) + bf16 output rounding (tests/test_gpu_kernels.py)
        assert torch.all(err <= 2 ** -6 * want.abs() + 6e-3), (tag, m, float(err.max()))
    return worst


def test_prefill_32k_prompt_in_8k_chunks_over_cached_prefix(cabi):
    """A fresh 32 768-token prompt computed as four 8 192-token chunks, each over the pages the earlier chunks
    left (the last one: 8 192 queries over a 24 576-token cached prefix), with a second, ragged sequence in the
    same launches. Hq=4 / Hkv=1 / head_dim 128 -> the tcgen05 kernel; block ids at the top of the 73 785-block
    pool, i.e. TMA row coordinates up to 73 785 * 16."""
    Hq, Hkv, D = 4, 1, 128
    g = torch.Generator().manual_seed(55)
    P, CH = 32768, 8192
    lens = [P, 3 * CH // 2 + 77]
    tables, bt = _tables(lens, g)
    kc, vc = _pool(Hkv, D)
    K = [torch.randn(L, Hkv, D, generator=g).to(torch.bfloat16) for L in lens]
    V = [_segment_values(L, Hkv, D, g, seg=512) for L in lens]   # O(1) outputs: errors cannot average out
    Q = [(2 * torch.randn(L, Hq, D, generator=g)).to(torch.bfloat16) for L in lens]
    btd = bt.cuda()
    worst = 0.0
    for c in range(P // CH):
        # chunk c of sequence 0; sequence 1 runs its own (shorter, ragged) chunks alongside while it has any
        spans = [(c * CH, CH)]
        a1 = c * 6000
        if a1 < lens[1]:
            spans.append((a1, min(6000, lens[1] - a1)))
        for s, (a, n) in enumerate(spans):     # the chunk's K/V rows reach the pages before attention reads them
            tbl = torch.tensor(tables[s], dtype=torch.long)
            pos = torch.arange(a, a + n)
            kc[tbl[pos // BS].cuda(), :, (pos % BS).cuda()] = K[s][a:a + n].cuda()
            vc[tbl[pos // BS].cuda(), :, (pos % BS).cuda()] = V[s][a:a + n].cuda()
        cu = [0]
        for _, n in spans:
            cu.append(cu[-1] + n)
        T = cu[-1]
        q = torch.cat([Q[s][a:a + n] for s, (a, n) in enumerate(spans)])
        out = torch.full((T, Hq * D), float("nan"), dtype=torch.bfloat16, device="cuda")
        cabi.paged_attn_prefill(q.cuda().view(T, Hq * D), kc, vc, btd[:len(spans)].contiguous(),
                                torch.tensor(cu, dtype=torch.int32).cuda(),
                                torch.tensor([a + n for a, n in spans], dtype=torch.int32).cuda(), out, len(spans),
                                max(n for _, n in spans), Hq, Hkv, D, BS, D ** -0.5)
        got = out.float().cpu().view(T, Hq, D)
        assert torch.isfinite(got).all()
        for s, (a, n) in enumerate(spans):
            wins = sorted({0, (n // 2 // 128) * 128 - 64 if n > 512 else 0, max(0, n - 128)})
            wins = [w for w in wins if w + 128 <= n] or [0]
            if n < 128:
                continue
            worst = max(worst, _check_windows(got[cu[s]:cu[s + 1]], Q[s][a:a + n], K[s].float(), V[s].float(), a, wins,
                                              (c, s)))
    print(f"\n[prefill 32768 in 8192-token chunks over cached pages] max-abs err {worst:.2e}")


def _shaped_keys(L, D, g, profile):
    """Keys whose scores against the common query direction follow `profile` (natural-log units, length L)."""
    u = torch.randn(D, generator=g)
    u = u / u.norm()
    k = 0.3 * torch.randn(L, D, generator=g)
    k = k - (k @ u)[:, None] * u          # noise orthogonal to u keeps the shaped component exact
    return u, k + profile[:, None] * u


def _segment_values(L, Hkv, D, g, seg=1024):
    """V rows that are (nearly) constant over segments of `seg` keys: the output is then the softmax MASS per
    segment times O(1) vectors, so a wrongly scaled accumulator shows up at full size instead of averaging out."""
    base = torch.randn((L + seg - 1) // seg, Hkv, D, generator=g)
    return (base[torch.arange(L) // seg] + 0.1 * torch.randn(L, Hkv, D, generator=g)).to(torch.bfloat16)


def test_prefill_lazy_rescale_branch_taken_and_not_taken(cabi):
    """attn_prefill.cu rescales O in TMEM only when a row's running maximum grew by more than 2^8 since the
    reference was last moved. Three score profiles over 16 384 keys drive that branch both ways, hundreds of
    tiles into the sequence:
      rising steps  - the maximum jumps by ~2^13 every 1 536 keys (rescale taken late and often; everything
                      accumulated before a jump must be scaled away exactly);
      slow rise     - the maximum creeps up by < 2^8 over the whole context (hardly ever rescaled after the
                      first tile: P grows beyond 1 against a stale reference, l and O accumulate against it);
      falling       - the largest scores come first (not rescaled; late P underflow towards 0).
    V is constant over 1 024-key segments, so the expected output weighs O(1) vectors by softmax mass."""
    Hq, Hkv, D = 4, 1, 128
    L, n = 16384, 1024
    g = torch.Generator().manual_seed(91)
    t = torch.arange(L, dtype=torch.float32)
    ln2 = 0.6931471805599453
    profiles = {
        "rising steps": 13 * ln2 * torch.floor(t / 1536) + 1.5 * torch.randn(L, generator=g),
        "slow rise": 4 * ln2 * t / L + 0.4 * torch.randn(L, generator=g),
        "falling": -20 * ln2 * t / L + 1.5 * torch.randn(L, generator=g),
    }
    lens = [L] * len(profiles)
    tables, bt = _tables(lens, g)
    kc, vc = _pool(Hkv, D)
    qs, Ks, Vs = [], [], []
    for tbl, (name, prof) in zip(tables, profiles.items()):
        # score = q.k / sqrt(D) with q = sqrt(D) * u (+ noise): the u-component of k IS the score
        u, k = _shaped_keys(L, D, g, prof)
        k = k[:, None, :].to(torch.bfloat16)
        v = _segment_values(L, Hkv, D, g)
        q = (D ** 0.5 * u[None, None, :] + 0.2 * torch.randn(n, Hq, D, generator=g)).to(torch.bfloat16)
        _scatter(kc, vc, tbl, k, v)
        qs.append(q)
        Ks.append(k.float())
        Vs.append(v.float())
    S = len(profiles)
    cu = [i * n for i in range(S + 1)]
    q_all = torch.cat(qs)
    out = torch.full((S * n, Hq * D), float("nan"), dtype=torch.bfloat16, device="cuda")
    cabi.paged_attn_prefill(q_all.cuda().view(S * n, Hq * D), kc, vc, bt.cuda(), torch.tensor(cu, dtype=torch.int32).cuda(),
                            torch.tensor(lens, dtype=torch.int32).cuda(), out, S, n, Hq, Hkv, D, BS, D ** -0.5)
    got = out.float().cpu().view(S * n, Hq, D)
    assert torch.isfinite(got).all()
    for i, name in enumerate(profiles):
        want = R.attention(qs[i].float(), Ks[i], Vs[i], L - n)
        err = (got[cu[i]:cu[i + 1]] - want).abs()
        print(f"\n[lazy rescale: {name}] max-abs err {float(err.max()):.2e}")
        assert torch.all(err <= 2 ** -6 * want.abs() + 6e-3), (name, float(err.max()))


def test_decode_attention_with_rising_scores_across_splits(cabi):
    """The split merge (attn_decode_reduce_kernel) with maxima that differ by thousands of e-folds between
    splits: partials of early splits must vanish exactly, not overflow or turn into NaN."""
    Hq, Hkv, D = 8, 1, 128
    g = torch.Generator().manual_seed(17)
    L = 32768
    t = torch.arange(L, dtype=torch.float32)
    profs = [60 * t / L + torch.randn(L, generator=g), -60 * t / L + torch.randn(L, generator=g),
             30 * torch.sin(t / 3000) + torch.randn(L, generator=g)]
    lens = [L] * len(profs)
    tables, bt = _tables(lens, g)
    kc, vc = _pool(Hkv, D)
    qs, Ks, Vs = [], [], []
    for tbl, prof in zip(tables, profs):
        u, k = _shaped_keys(L, D, g, prof)
        k = k[:, None, :].to(torch.bfloat16)
        v = _segment_values(L, Hkv, D, g, seg=256)
        _scatter(kc, vc, tbl, k, v)
        qs.append((D ** 0.5 * u[None, :] + 0.2 * torch.randn(Hq, D, generator=g)).to(torch.bfloat16))
        Ks.append(k.float())
        Vs.append(v.float())
    B = len(profs)
    q = torch.stack(qs)
    for splits in (16, 32):
        out = torch.full((B, Hq * D), float("nan"), dtype=torch.bfloat16, device="cuda")
        ws = torch.empty(cabi.decode_workspace_bytes(B, Hq, D, splits) // 4, device="cuda")
        cabi.paged_attn_decode(q.cuda().view(B, Hq * D), kc, vc, bt.cuda(), torch.tensor(lens, dtype=torch.int32).cuda(),
                               out, ws, B, Hq, Hkv, D, BS, splits, D ** -0.5)
        got = out.float().cpu().view(B, Hq, D)
        assert torch.isfinite(got).all()
        for b in range(B):
            want = R.attention(q[b].float()[None], Ks[b], Vs[b], L - 1)[0]
            assert torch.all((got[b] - want).abs() <= 2 ** -7 * want.abs() + 2e-3), (splits, b)
