"""Per-kernel parity: every C-ABI entry point against the CPU oracle on identical (bf16-rounded)
inputs. Integer work (page scatter, arg-max) is bit-exact; floating point uses the tolerance
stated beside each check."""
import math

import numpy as np
import pytest
import torch

from oracle import llama_ref as R
from oracle import sampler_ref as SR

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cabi():
    from atx import _cabi

    _cabi.load()
    return _cabi


def _bf16(t):
    return t.to(torch.bfloat16)


# ---------------------------------------------------------------------------------------- GEMM
GEMM_CASES = [
    # T, N, K, mode, splits
    (1, 128, 64, 0, 1),
    (6, 256, 512, 0, 1),
    (6, 384, 4096, 1, 4),
    (16, 128, 128, 2, 1),
    (17, 256, 1024, 0, 1),
    (33, 512, 256, 1, 2),
    (100, 256, 512, 2, 1),
    (130, 384, 1024, 0, 1),
    (300, 256, 2048, 1, 1),
    (6, 200, 192, 0, 1),       # N not a multiple of 128: TMA zero fill + guarded stores
    (6, 6144, 4096, 1, 0),     # 8B qkv shape, automatic split-K
    (512, 1024, 4096, 2, 1),   # prefill-like tile (TN = 256, two token tiles)
]


@pytest.mark.parametrize("T,N,K,mode,splits", GEMM_CASES)
def test_gemm_matches_fp32_reference(cabi, T, N, K, mode, splits):
    g = torch.Generator().manual_seed(T * 131 + N * 7 + K)
    x = _bf16(torch.randn(T, K, generator=g))
    w = _bf16(torch.randn(N, K, generator=g) * 0.05)
    ref = x.float() @ w.float().T  # exact products of bf16 values, fp32/64-safe reference
    ref64 = (x.double() @ w.double().T)
    t_cap = max(T, 8)
    xb = torch.zeros(t_cap, K, dtype=torch.bfloat16, device="cuda")
    xb[:T] = x.cuda()
    plan = cabi.GemmPlan(w.cuda().contiguous(), xb)
    if mode == 2:
        out = torch.full((T, N), 7.0, dtype=torch.bfloat16, device="cuda")
    elif mode == 1:
        base = torch.randn(T, N, generator=g)
        out = base.cuda().clone()
    else:
        out = torch.full((T, N), 7.0, dtype=torch.float32, device="cuda")
    plan.run(out, T, mode, splits)
    torch.cuda.synchronize()
    got = out.float().cpu().double()
    want = ref64 + (base.double() if mode == 1 else 0)
    # fp32 accumulation over K terms of magnitude ~|x||w|: relative 1e-5 of the row scale + bf16 output rounding
    scale = (x.double().abs() @ w.double().abs().T)
    tol = 2e-6 * scale + (2 ** -8 * want.abs() if mode == 2 else 0) + 1e-6
    assert torch.all((got - want).abs() <= tol), float(((got - want).abs() - tol).max())
    assert ref.shape == got.shape


@pytest.mark.parametrize("T,N,K,splits", [(6, 4096, 4096, 0), (6, 6144, 4096, 6), (32, 4096, 14336, 0), (64, 512, 1024, 8),
                                          (1, 128, 64, 0)])
def test_gemm_ordered_split_k_is_bit_reproducible(cabi, T, N, K, splits):
    """ATX_FLAG_ORDERED: the K-splits of an ADD_F32 are summed in a fixed order inside a cluster and one CTA adds
    each output element, so repeated runs agree bit for bit (plain split-K lets every split race its red.add);
    the sum itself is checked against the fp64 reference like every other GEMM case."""
    g = torch.Generator().manual_seed(T + N + K)
    x = _bf16(torch.randn(T, K, generator=g))
    w = _bf16(torch.randn(N, K, generator=g) * 0.05)
    base = torch.randn(T, N, generator=g)
    xb = torch.zeros(max(T, 8), K, dtype=torch.bfloat16, device="cuda")
    xb[:T] = x.cuda()
    plan = cabi.GemmPlan(w.cuda().contiguous(), xb)
    runs = []
    for _ in range(4):
        out = base.cuda().clone()
        plan.run(out, T, cabi.GEMM_ADD_F32, splits, flags=cabi.FLAG_ORDERED)
        torch.cuda.synchronize()
        runs.append(out.cpu())
    assert all(torch.equal(runs[0], r) for r in runs[1:])
    want = x.double() @ w.double().T + base.double()
    scale = x.double().abs() @ w.double().abs().T
    assert torch.all((runs[0].double() - want).abs() <= 2e-6 * scale + 1e-6)


@pytest.mark.parametrize("T,I,K", [(6, 128, 256), (16, 1024, 512), (40, 192, 128), (300, 256, 1024)])
def test_gemm_fused_swiglu_epilogue(cabi, T, I, K):
    """ATX_GEMM_SWIGLU_BF16: interleaved gate/up rows, act = silu(x Wg^T) * (x Wu^T) in bf16."""
    from atx.weights import interleave_gate_up

    g = torch.Generator().manual_seed(T + I + K)
    x = _bf16(torch.randn(T, K, generator=g))
    wg = _bf16(torch.randn(I, K, generator=g) * 0.08)
    wu = _bf16(torch.randn(I, K, generator=g) * 0.08)
    xb = torch.zeros(max(T, 8), K, dtype=torch.bfloat16, device="cuda")
    xb[:T] = x.cuda()
    plan = cabi.GemmPlan(interleave_gate_up(wg, wu).cuda().contiguous(), xb)
    out = torch.full((T, I), 7.0, dtype=torch.bfloat16, device="cuda")
    plan.run(out, T, cabi.GEMM_SWIGLU_BF16, 1)
    want = R.swiglu(x.float() @ wg.float().T, x.float() @ wu.float().T)
    err = (out.float().cpu() - want).abs()
    assert torch.all(err <= 2 ** -7 * want.abs() + 1e-3), float(err.max())
    with pytest.raises(cabi.AtxError, match="split-K"):
        plan.run(out, T, cabi.GEMM_SWIGLU_BF16, 3)   # cluster split-K takes 2, 4 or 8 only
    if T > 64:
        with pytest.raises(cabi.AtxError, match="split-K"):
            plan.run(out, T, cabi.GEMM_SWIGLU_BF16, 2)


@pytest.mark.parametrize("T,I,K,splits", [(1, 128, 512, 2), (6, 448, 1024, 4), (16, 192, 2048, 8), (17, 256, 512, 8),
                                          (32, 3584, 8192, 0), (33, 320, 1024, 2), (64, 128, 512, 4)])
def test_gemm_swiglu_cluster_split_k(cabi, T, I, K, splits):
    """SwiGLU with split-K: the K-splits of a tile are one thread-block cluster, partial sums meet in the
    leader's TMEM through distributed shared memory. Same result as the unsplit kernel up to fp32 summation
    order (compared in bf16), NaN-poisoned destination fully overwritten, repeatable bit for bit."""
    from atx.weights import interleave_gate_up

    g = torch.Generator().manual_seed(11 * T + I + K)
    x = _bf16(torch.randn(T, K, generator=g))
    wg = _bf16(torch.randn(I, K, generator=g) * 0.05)
    wu = _bf16(torch.randn(I, K, generator=g) * 0.05)
    xb = torch.full((max(T, 8) + 3, K), float("nan"), dtype=torch.bfloat16, device="cuda")
    xb[:T] = x.cuda()
    plan = cabi.GemmPlan(interleave_gate_up(wg, wu).cuda().contiguous(), xb)
    outs = []
    for _ in range(2):
        out = torch.full((T, I), float("nan"), dtype=torch.bfloat16, device="cuda")
        plan.run(out, T, cabi.GEMM_SWIGLU_BF16, splits)
        outs.append(out.clone())
    assert torch.equal(outs[0], outs[1])
    want = R.swiglu(x.float() @ wg.float().T, x.float() @ wu.float().T)
    err = (outs[0].float().cpu() - want).abs()
    assert torch.all(err <= 2 ** -7 * want.abs() + 1e-3), float(err.max())
    one = torch.empty_like(outs[0])
    plan.run(one, T, cabi.GEMM_SWIGLU_BF16, 1)
    d = (one.float() - outs[0].float()).abs()
    assert torch.all(d <= 2 ** -7 * one.float().abs() + 1e-3), float(d.max())


def test_gemm_rows_beyond_T_do_not_leak(cabi):
    """Rows of the X buffer past T (stale data, even NaN) must not influence rows < T."""
    K, N, T = 256, 256, 5
    g = torch.Generator().manual_seed(3)
    xb = _bf16(torch.randn(16, K, generator=g)).cuda()
    xb[T:] = float("nan")
    w = _bf16(torch.randn(N, K, generator=g)).cuda()
    out = torch.zeros(16, N, device="cuda")
    cabi.GemmPlan(w, xb).run(out, T, 0)
    torch.cuda.synchronize()
    want = xb[:T].float().cpu() @ w.float().cpu().T
    assert torch.allclose(out[:T].cpu(), want, rtol=1e-4, atol=1e-3)
    assert torch.all(out[T:] == 0)


def test_gemm_rejects_bad_arguments(cabi):
    w = torch.zeros(128, 96, dtype=torch.bfloat16, device="cuda")
    x = torch.zeros(8, 96, dtype=torch.bfloat16, device="cuda")
    with pytest.raises(cabi.AtxError, match="multiple of 64"):
        cabi.GemmPlan(w, x)
    w = torch.zeros(128, 128, dtype=torch.bfloat16, device="cuda")
    x = torch.zeros(8, 128, dtype=torch.bfloat16, device="cuda")
    plan = cabi.GemmPlan(w, x)
    out = torch.zeros(8, 128, device="cuda")
    with pytest.raises(cabi.AtxError, match="outside"):
        plan.run(torch.zeros(16, 128, device="cuda"), 9, 0)
    with pytest.raises(cabi.AtxError, match="split-K"):
        plan.run(out, 4, 0, splits=2)


# ---------------------------------------------------------------------------------- pointwise
def test_embed_gather_exact(cabi):
    g = torch.Generator().manual_seed(5)
    table = _bf16(torch.randn(1000, 512, generator=g))
    ids = torch.randint(0, 1000, (37,), generator=g, dtype=torch.int32)
    out = torch.empty(37, 512, device="cuda")
    cabi.embed_gather(ids.cuda(), table.cuda(), out, 37)
    assert torch.equal(out.cpu(), table.float()[ids.long()])


@pytest.mark.parametrize("T,H", [(1, 512), (6, 4096), (33, 2048), (5, 8192)])
def test_rmsnorm(cabi, T, H):
    g = torch.Generator().manual_seed(T + H)
    x = torch.randn(T, H, generator=g) * 3
    w = _bf16(1 + 0.1 * torch.randn(H, generator=g))
    out = torch.empty(T, H, dtype=torch.bfloat16, device="cuda")
    zero = torch.full((1000, 4), 5.0, device="cuda")
    cabi.rmsnorm(x.cuda(), w.cuda(), out, T, 1e-5, zero=zero)
    want = R.rms_norm(x, w.float(), 1e-5)
    # one bf16 rounding of an O(1) value computed in fp32
    assert torch.all((out.float().cpu() - want).abs() <= 2 ** -8 * want.abs() + 1e-6)
    assert torch.all(zero == 0)


@pytest.mark.parametrize("rows", [False, True])
def test_rmsnorm_residual_add_fp32_and_bf16(cabi, rows):
    """Tensor-parallel residual: x += add before the norm. fp32 add (decode-sized steps: split-K partials, handed back
    cleared) and bf16 add (ATX_FLAG_ADD_BF16: prefill-sized steps all-reduce in bf16; left untouched). With a row
    gather nothing is written back."""
    g = torch.Generator().manual_seed(31)
    T, H = 12, 4096
    x0 = torch.randn(T, H, generator=g) * 2
    a0 = torch.randn(T, H, generator=g)
    w = _bf16(1 + 0.1 * torch.randn(H, generator=g))
    idx = torch.tensor([11, 0, 5], dtype=torch.int32) if rows else None
    n = 3 if rows else T
    sel = idx.long() if rows else torch.arange(T)
    for flag, add0 in ((0, a0), (cabi.FLAG_ADD_BF16, _bf16(a0))):
        x, add = x0.cuda(), add0.cuda()
        out = torch.empty(n, H, dtype=torch.bfloat16, device="cuda")
        cabi.rmsnorm(x, w.cuda(), out, n, 1e-5, rows=idx.cuda() if rows else None, add=add, flags=flag)
        summed = x0[sel] + add0.float()[sel]
        want = R.rms_norm(summed, w.float(), 1e-5)
        assert torch.all((out.float().cpu() - want).abs() <= 2 ** -8 * want.abs() + 1e-6)
        if rows:
            assert torch.equal(x.cpu(), x0) and torch.equal(add.cpu(), add0)
        else:
            assert torch.equal(x.cpu(), summed)                                  # residual written back
            assert torch.equal(add.cpu(), add0 if flag else torch.zeros(T, H))   # fp32 buffer handed back cleared


def test_rmsnorm_row_gather(cabi):
    g = torch.Generator().manual_seed(9)
    x = torch.randn(20, 512, generator=g)
    w = _bf16(torch.ones(512))
    rows = torch.tensor([19, 3, 7], dtype=torch.int32)
    out = torch.empty(3, 512, dtype=torch.bfloat16, device="cuda")
    cabi.rmsnorm(x.cuda(), w.cuda(), out, 3, 1e-5, rows=rows.cuda())
    want = R.rms_norm(x[rows.long()], w.float(), 1e-5)
    assert torch.all((out.float().cpu() - want).abs() <= 2 ** -8 * want.abs() + 1e-6)


@pytest.mark.parametrize("shape_name", ["tiny", "8b"])
def test_rope_kv_write_bit_exact(cabi, shape_name):
    """q (bf16) and the K/V pages must equal the oracle's rope + bf16 rounding exactly, and the
    scatter must hit exactly the slots of the block table (integer indexing: bit-exact)."""
    s = R.shape_by_name(shape_name)
    Hq, Hkv, D, bs = s.num_attention_heads, s.num_key_value_heads, s.head_dim, 16
    T, nblocks = 23, 12
    g = torch.Generator().manual_seed(17)
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, generator=g)
    positions = torch.tensor([0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 100, 101, 102, 8191, 8192, 8193, 20000, 20001, 30000,
                              30001, 30002, 32767, 5], dtype=torch.int32)
    table = [7, 2, 9]
    slots = torch.tensor(R.slot_mapping(table, 10, T - 1, bs) + [-1], dtype=torch.int32)
    cs = R.cos_sin_table(s, 32768)
    kc = torch.zeros(nblocks, Hkv, bs, D, dtype=torch.bfloat16, device="cuda")
    vc = torch.zeros_like(kc)
    qo = torch.empty(T, Hq * D, dtype=torch.bfloat16, device="cuda")
    cabi.rope_kv_write(qkv.cuda(), positions.cuda(), slots.cuda(), cs.cuda(), qo, kc, vc, T, Hq, Hkv, D, bs)
    # The kernel looks cos/sin up in the table, so the bit-exact statement is: kernel == oracle arithmetic on the
    # table's rows. That the table equals a direct evaluation at the same positions is a property of the host's
    # libm, checked separately (a vectorised cos may pick its slow path per SIMD vector, so two evaluations of
    # the same angle can differ in the last bit depending on their neighbours; seen once on a GPU box).
    half = D // 2
    rows_ = cs[positions.long()]
    cos = torch.cat((rows_[:, :half], rows_[:, :half]), dim=-1)
    sin = torch.cat((rows_[:, half:], rows_[:, half:]), dim=-1)
    cos_d, sin_d = R.rope_cos_sin(positions, R.llama3_inv_freq(s))
    # (host libm property, a few ulp whichever vector / scalar path each evaluation took; failed once at 2^-23 on a
    # GPU box when run after other tests, passed alone on the same box)
    dc, ds = float((cos - cos_d).abs().max()), float((sin - sin_d).abs().max())
    assert dc <= 2 ** -21 and ds <= 2 ** -21, f"cos/sin table vs direct evaluation differ by {dc:.3e} / {ds:.3e}"
    q = qkv[:, : Hq * D].view(T, Hq, D)
    k = qkv[:, Hq * D: (Hq + Hkv) * D].view(T, Hkv, D)
    v = qkv[:, (Hq + Hkv) * D:].view(T, Hkv, D)
    q_ref = R.apply_rope(q, cos, sin).to(torch.bfloat16)
    k_ref = R.apply_rope(k, cos, sin).to(torch.bfloat16)
    got_q = qo.cpu().view(T, Hq, D)
    if not torch.equal(got_q, q_ref):   # say exactly what differs: position, head, dim, values, table vs direct cos
        bad = (got_q != q_ref).nonzero().tolist()
        t, h, i = bad[0]
        tbl = cs[int(positions[t])]
        raise AssertionError(
            f"{len(bad)} q elements differ, first at token {t} (pos {int(positions[t])}) head {h} dim {i}: "
            f"kernel {got_q[t, h, i].item()} oracle {q_ref[t, h, i].item()}; x={q[t, h, i].item()!r} "
            f"partner={q[t, h, (i + half) % D].item()!r}; table cos/sin {tbl[i % half].item()!r}/{tbl[half + i % half].item()!r} "
            f"direct {cos_d[t, i].item()!r}/{sin_d[t, i].item()!r}; tokens affected {sorted({b[0] for b in bad})}")
    kc_ref = torch.zeros(nblocks, Hkv, bs, D, dtype=torch.bfloat16)
    vc_ref = torch.zeros_like(kc_ref)
    for t in range(T):
        sl = int(slots[t])
        if sl < 0:
            continue
        kc_ref[sl // bs, :, sl % bs] = k_ref[t]
        vc_ref[sl // bs, :, sl % bs] = v[t].to(torch.bfloat16)
    assert torch.equal(kc.cpu(), kc_ref), f"{int((kc.cpu() != kc_ref).sum())} K-page elements differ"
    assert torch.equal(vc.cpu(), vc_ref), f"{int((vc.cpu() != vc_ref).sum())} V-page elements differ"


def test_silu_mul(cabi):
    g = torch.Generator().manual_seed(23)
    T, I = 7, 1024
    gu = torch.randn(T, 2 * I, generator=g) * 4
    act = torch.empty(T, I, dtype=torch.bfloat16, device="cuda")
    cabi.silu_mul(gu.cuda(), act, T, I)
    want = R.swiglu(gu[:, :I], gu[:, I:])
    assert torch.all((act.float().cpu() - want).abs() <= 2 ** -8 * want.abs() + 1e-6)


# ---------------------------------------------------------------------------------- attention
def _paged_case(seed, lens, Hq, Hkv, D, bs=16, extra_blocks=5):
    """Random K/V for every sequence scattered into shuffled pages (+ NaN in unused rows)."""
    g = torch.Generator().manual_seed(seed)
    need = [(L + bs - 1) // bs for L in lens]
    nblocks = sum(need) + extra_blocks
    perm = torch.randperm(nblocks, generator=g).tolist()
    kc = torch.full((nblocks, Hkv, bs, D), float("nan"), dtype=torch.bfloat16)
    vc = torch.full_like(kc, float("nan"))
    tables, ks, vs = [], [], []
    max_blocks = max(need) + 2
    for i, L in enumerate(lens):
        tbl = [perm.pop() for _ in range(need[i])]
        k = _bf16(torch.randn(L, Hkv, D, generator=g))
        v = _bf16(torch.randn(L, Hkv, D, generator=g))
        for p in range(L):
            kc[tbl[p // bs], :, p % bs] = k[p]
            vc[tbl[p // bs], :, p % bs] = v[p]
        tables.append(tbl + [0] * (max_blocks - len(tbl)))
        ks.append(k.float())
        vs.append(v.float())
    return kc, vc, torch.tensor(tables, dtype=torch.int32), ks, vs


@pytest.mark.parametrize("Hq,Hkv,D", [(32, 8, 128), (8, 2, 64), (8, 1, 128), (4, 4, 64)])
@pytest.mark.parametrize("splits", [1, 3, 8])
def test_paged_decode_attention(cabi, Hq, Hkv, D, splits):
    lens = [1, 15, 16, 17, 100, 257, 1153]
    kc, vc, bt, ks, vs = _paged_case(31 + Hq + D + splits, lens, Hq, Hkv, D)
    B = len(lens)
    g = torch.Generator().manual_seed(77)
    q = _bf16(torch.randn(B, Hq, D, generator=g))
    out = torch.empty(B, Hq * D, dtype=torch.bfloat16, device="cuda")
    ws = torch.empty(max(1, cabi.decode_workspace_bytes(B, Hq, D, splits) // 4), device="cuda")
    cabi.paged_attn_decode(q.cuda().view(B, Hq * D), kc.cuda(), vc.cuda(), bt.cuda(),
                           torch.tensor(lens, dtype=torch.int32).cuda(), out, ws, B, Hq, Hkv, D, 16, splits,
                           D ** -0.5)
    got = out.float().cpu().view(B, Hq, D)
    for b, L in enumerate(lens):
        want = R.attention(q[b].float()[None], ks[b], vs[b], L - 1)[0]
        # fp32 softmax/accumulate, one bf16 rounding of an O(1) output
        assert torch.all((got[b] - want).abs() <= 2 ** -7 * want.abs() + 2e-3), (b, L)


def test_paged_decode_zero_length_rows(cabi):
    """Padded batch entries (seq_len 0) must produce zeros, not NaN."""
    kc, vc, bt, ks, vs = _paged_case(5, [40, 40], 8, 2, 64)
    q = _bf16(torch.randn(2, 8 * 64))
    out = torch.full((2, 8 * 64), 9.0, dtype=torch.bfloat16, device="cuda")
    ws = torch.empty(cabi.decode_workspace_bytes(2, 8, 64, 4) // 4, device="cuda")
    cabi.paged_attn_decode(q.cuda(), kc.cuda(), vc.cuda(), bt.cuda(), torch.tensor([40, 0], dtype=torch.int32).cuda(),
                           out, ws, 2, 8, 2, 64, 16, 4, 0.125)
    assert torch.all(out[1] == 0) and torch.isfinite(out.float()).all()


@pytest.mark.parametrize("Hq,Hkv,D", [(32, 8, 128), (8, 2, 64)])
def test_paged_prefill_attention(cabi, Hq, Hkv, D):
    # (context length incl. chunk, chunk length): fresh prompts, chunked continuation, prefix hit, 1-token chunk
    cases = [(64, 64), (200, 200), (333, 77), (1024, 256), (17, 1), (130, 130)]
    lens = [c for c, _ in cases]
    kc, vc, bt, ks, vs = _paged_case(41 + D, lens, Hq, Hkv, D)
    g = torch.Generator().manual_seed(78)
    nq = [n for _, n in cases]
    cu = [0]
    for n in nq:
        cu.append(cu[-1] + n)
    T = cu[-1]
    q = _bf16(torch.randn(T, Hq, D, generator=g))
    out = torch.full((T, Hq * D), float("nan"), dtype=torch.bfloat16, device="cuda")
    cabi.paged_attn_prefill(q.cuda().view(T, Hq * D), kc.cuda(), vc.cuda(), bt.cuda(),
                            torch.tensor(cu, dtype=torch.int32).cuda(), torch.tensor(lens, dtype=torch.int32).cuda(),
                            out, len(cases), max(nq), Hq, Hkv, D, 16, D ** -0.5)
    got = out.float().cpu().view(T, Hq, D)
    for i, (ctx, n) in enumerate(cases):
        want = R.attention(q[cu[i]:cu[i + 1]].float(), ks[i], vs[i], ctx - n)
        err = (got[cu[i]:cu[i + 1]] - want).abs()
        # P is rounded to bf16 before P.V (2^-9 relative per term) + bf16 output rounding
        assert torch.all(err <= 2 ** -6 * want.abs() + 6e-3), (i, float(err.max()))


@pytest.mark.parametrize("Hq,Hkv,D", [(8, 1, 128), (4, 1, 64)])
def test_paged_prefill_attention_tile_edges(cabi, Hq, Hkv, D):
    """Sizes around the kernel's tiling: 128-row query tiles in pairs (one tile only, a second tile with one row),
    128-key K/V tiles handed over in 64-key halves (a half that is fully masked or past the context, contexts
    ending on and next to page / half / tile boundaries, tile B seeing one key tile more than tile A), GQA groups
    of 8 (70B) and 4, chunk offsets on and off a tile boundary."""
    cases = [(1, 1), (63, 63), (64, 1), (65, 65), (127, 127), (128, 128), (129, 129), (192, 64), (255, 255),
             (256, 256), (257, 257), (300, 44), (513, 385), (1000, 1000), (70, 3), (384, 128), (385, 129)]
    lens = [c for c, _ in cases]
    kc, vc, bt, ks, vs = _paged_case(7 + D, lens, Hq, Hkv, D)
    g = torch.Generator().manual_seed(79)
    nq = [n for _, n in cases]
    cu = [0]
    for n in nq:
        cu.append(cu[-1] + n)
    T = cu[-1]
    q = _bf16(torch.randn(T, Hq, D, generator=g))
    out = torch.full((T, Hq * D), float("nan"), dtype=torch.bfloat16, device="cuda")
    cabi.paged_attn_prefill(q.cuda().view(T, Hq * D), kc.cuda(), vc.cuda(), bt.cuda(),
                            torch.tensor(cu, dtype=torch.int32).cuda(), torch.tensor(lens, dtype=torch.int32).cuda(),
                            out, len(cases), max(nq), Hq, Hkv, D, 16, D ** -0.5)
    got = out.float().cpu().view(T, Hq, D)
    assert torch.isfinite(got).all()
    for i, (ctx, n) in enumerate(cases):
        want = R.attention(q[cu[i]:cu[i + 1]].float(), ks[i], vs[i], ctx - n)
        err = (got[cu[i]:cu[i + 1]] - want).abs()
        assert torch.all(err <= 2 ** -6 * want.abs() + 6e-3), ((ctx, n), float(err.max()))


# ------------------------------------------------------------------------------------- sampler
def test_argmax_exact_and_tie_break(cabi):
    g = torch.Generator().manual_seed(1)
    V = 128256
    logits = torch.randn(5, V, generator=g)
    logits[2, 777] = logits[2, 99999] = 50.0  # tie: lowest index wins (torch.argmax)
    logits[3, V - 1] = 60.0
    out = torch.empty(5, dtype=torch.int32, device="cuda")
    cabi.sample(logits.cuda(), out, 5, V)
    assert out.cpu().tolist() == torch.argmax(logits, dim=-1).tolist()
    assert out[2].item() == 777


def test_temperature_sampling_matches_oracle_stream(cabi):
    V, B = 4096, 16
    g = torch.Generator().manual_seed(2)
    logits = torch.randn(B, V, generator=g) * 2
    temps = torch.full((B,), 0.7)
    temps[0] = 0.0
    seeds = torch.arange(B, dtype=torch.int64) * 7919 + (1 << 40)
    steps = torch.arange(B, dtype=torch.int32)
    out = torch.empty(B, dtype=torch.int32, device="cuda")
    cabi.sample(logits.cuda(), out, B, V, temps.cuda(), None, seeds.cuda(), steps.cuda())
    ids = out.cpu().tolist()
    assert ids[0] == int(torch.argmax(logits[0]))
    for b in range(1, B):
        sc = SR.sample_scores(logits[b].numpy(), 0.7, int(seeds[b]), int(steps[b]))
        # logf on the device may differ from numpy by an ulp: the chosen id must be the oracle's
        # arg-max up to 1e-4 in score
        assert sc[ids[b]] >= sc.max() - 1e-4, b


@pytest.mark.parametrize("world", [2, 8])
def test_vocabulary_parallel_sampling_equals_the_unsharded_sampler(cabi, world):
    """atx_sample_shard on each vocabulary shard + atx_sample_merge over the shards' (score, index) pairs draws
    exactly the token atx_sample draws from the whole row: greedy rows (with a tie that straddles two shards: the
    lowest index wins) and temperature rows (Gumbel noise keyed by the global index)."""
    V, B = 128256, 12
    Vl = V // world
    g = torch.Generator().manual_seed(31 + world)
    logits = torch.randn(B, V, generator=g) * 2
    logits[1, Vl - 1] = logits[1, Vl] = 40.0           # tie across the boundary of shards 0 and 1
    logits[2, V - 1] = 45.0
    temps = torch.tensor([0.0, 0.0, 0.0] + [0.2, 0.7, 1.0] * 3)
    seeds = torch.arange(B, dtype=torch.int64) * 104729 + (1 << 41)
    steps = torch.arange(B, dtype=torch.int32) * 3
    d, temps, seeds, steps = logits.cuda(), temps.cuda(), seeds.cuda(), steps.cuda()
    want = torch.empty(B, dtype=torch.int32, device="cuda")
    cabi.sample(d, want, B, V, temps, None, seeds, steps)
    pairs = torch.zeros(world, B, 2, device="cuda")
    for r in range(world):
        shard = d[:, r * Vl:(r + 1) * Vl]                # a strided view: ldl = V, as in the engine
        cabi._check(cabi.load().atx_sample_shard(shard.data_ptr(), V, r * Vl, temps.data_ptr(), seeds.data_ptr(),
                                                  steps.data_ptr(), pairs[r].data_ptr(), B, Vl, 0,
                                                  torch.cuda.current_stream().cuda_stream))
    got = torch.empty(B, dtype=torch.int32, device="cuda")
    cabi.sample_merge(pairs, world, B, got)
    assert got.cpu().tolist() == want.cpu().tolist()
    assert got[1].item() == Vl - 1 and got[2].item() == V - 1


def test_temperature_sampling_distribution(cabi):
    """Chi-square style check of the draw against softmax(l / T) on a small vocabulary."""
    V, N = 8, 4096
    logits = torch.tensor([[2.0, 1.0, 0.5, 0.0, -0.5, -1.0, -3.0, 1.5]]).repeat(N, 1)
    temps = torch.full((N,), 0.8)
    seeds = torch.arange(N, dtype=torch.int64) + 12345
    steps = torch.zeros(N, dtype=torch.int32)
    out = torch.empty(N, dtype=torch.int32, device="cuda")
    cabi.sample(logits.cuda(), out, N, V, temps.cuda(), None, seeds.cuda(), steps.cuda())
    freq = np.bincount(out.cpu().numpy(), minlength=V) / N
    p = torch.softmax(logits[0] / 0.8, -1).numpy()
    assert np.all(np.abs(freq - p) < 4 * np.sqrt(p * (1 - p) / N) + 2e-3), (freq, p)


def test_top_p_restricts_to_nucleus(cabi):
    V, N = 64, 512
    g = torch.Generator().manual_seed(4)
    row = torch.randn(V, generator=g) * 3
    logits = row.repeat(N, 1)
    T, P = 1.0, 0.6
    keep = SR.nucleus_mask(row.numpy(), T, P)
    out = torch.empty(N, dtype=torch.int32, device="cuda")
    cabi.sample(logits.cuda(), out, N, V, torch.full((N,), T).cuda(), torch.full((N,), P).cuda(),
                (torch.arange(N, dtype=torch.int64) + 99).cuda(), torch.zeros(N, dtype=torch.int32).cuda())
    ids = out.cpu().numpy()
    assert keep[ids].all()
    assert len(set(ids.tolist())) == keep.sum() or keep.sum() > 8  # every nucleus token is reachable


@pytest.mark.parametrize("T,H,world,use_rows", [(1, 512, 2, False), (5, 4096, 2, False), (32, 8192, 8, False),
                                                (7, 6144, 4, True), (13, 8192, 8, True), (3, 16384, 2, False)])
def test_reduce_scatter_rmsnorm_all_gather_one_gpu_fake_peers(cabi, T, H, world, use_rows):
    """Two-shot form of the fused all-reduce on ONE GPU: the kernel is run once as every rank in turn (each with
    its own residual copy, state and flag array; the other ranks' announcements pre-set by the host). Afterwards
    every "rank's" activation buffer must hold ALL normalised rows (each produced by the rank owning that
    residual row and pushed everywhere), and each residual copy must be updated exactly on its own rows."""
    import ctypes as C

    lib = cabi.load()
    g = torch.Generator().manual_seed(T * 17 + H + world)
    R_rows = T + 3
    x0 = torch.randn(R_rows, H, generator=g)
    parts = [torch.randn(R_rows, H, generator=g) * 0.5 for _ in range(world)]
    w = _bf16(1 + 0.1 * torch.randn(H, generator=g))
    rows = torch.tensor([(3 * i + 1) % R_rows for i in range(T)], dtype=torch.int32) if use_rows else None
    xs = [x0.cuda() for _ in range(world)]
    dparts = [p.cuda() for p in parts]
    outs = [torch.full((T, H), float("nan"), dtype=torch.bfloat16, device="cuda") for _ in range(world)]
    flags = torch.full((world, 64), 1 << 30, dtype=torch.int32, device="cuda")
    states = [torch.zeros(2, dtype=torch.int32, device="cuda") for _ in range(world)]
    clear = torch.full((R_rows, H), 3.0, device="cuda")
    zero = torch.full((4096,), 9.0, device="cuda")
    wd = w.cuda()
    vp = C.c_void_p
    part_ptrs = (vp * world)(*[p.data_ptr() for p in dparts])
    flag_ptrs = (vp * world)(*[flags[r].data_ptr() for r in range(world)])
    out_ptrs = (vp * world)(*[o.data_ptr() for o in outs])
    rows_d = rows.cuda() if rows is not None else None
    st = torch.cuda.current_stream().cuda_stream
    for rk in range(world):
        # slots other ranks read from rank rk start below the epoch, as on a real run
        flags[:, rk] = 0
        flags[:, world + rk] = 0
        rc = lib.atx_reduce_scatter_rmsnorm_all_gather(
            xs[rk].data_ptr(), part_ptrs, flag_ptrs, out_ptrs, clear.data_ptr(), R_rows, states[rk].data_ptr(), rk, world,
            rows_d.data_ptr() if rows_d is not None else None, wd.data_ptr(), T, H, 1e-5, zero.data_ptr(),
            zero.numel() * 4, 0, st)
        assert rc == 0, lib.atx_last_error().decode()
        torch.cuda.synchronize()
        assert torch.all(clear == 0) and torch.all(zero == 0)
        assert int(states[rk][0]) == 1 and int(states[rk][1]) == 0
        assert torch.all(flags[:, rk] == 1) and torch.all(flags[:, world + rk] == 1)   # both barriers announced
        clear.fill_(3.0)
        zero.fill_(9.0)
    idx = rows.long() if rows is not None else torch.arange(T)
    summed = x0[idx].clone()
    for p in parts:
        summed += p[idx]
    want = R.rms_norm(summed, w.float(), 1e-5)
    for rk in range(world):
        got = outs[rk].float().cpu()
        assert torch.all((got - want).abs() <= 2 ** -8 * want.abs() + 1e-6), (rk, float((got - want).abs().max()))
        assert torch.equal(outs[rk], outs[0])
        want_x = x0.clone()
        if rows is None:
            mine = [t for t in range(T) if t % world == rk]
            want_x[mine] = summed[mine]
        assert torch.equal(xs[rk].cpu(), want_x)


@pytest.mark.parametrize("T,H,world,use_rows", [(1, 512, 2, False), (5, 4096, 2, False), (32, 8192, 8, False),
                                                (7, 6144, 4, True), (13, 8192, 8, True), (3, 16384, 2, False)])
def test_allreduce_ll_one_gpu_fake_peers(cabi, T, H, world, use_rows):
    """Low-latency (push, data-is-the-flag) all-reduce + add + RMSNorm on ONE GPU, run as rank `rk` with the host
    playing the other ranks: their partial rows are pre-written into rk's receive slots and the rows rk does not
    own are pre-written into rk's activation buffer, so every spin in the kernel finds its data. Checks what rk
    pushed to each owner, the rows it produced in EVERY rank's activation buffer, its residual rows, the
    re-armed slots and the housekeeping regions, for every rk."""
    import ctypes as C

    lib = cabi.load()
    g = torch.Generator().manual_seed(T * 29 + H + world)
    R_rows, T_cap = T + 3, T + 2
    x0 = torch.randn(R_rows, H, generator=g)
    parts = [torch.randn(R_rows, H, generator=g) * 0.5 for _ in range(world)]
    for p in parts:
        p[0, :8] = -0.0            # a payload equal to the sentinel must be delivered as +0.0, not wait forever
    w = _bf16(1 + 0.1 * torch.randn(H, generator=g))
    rows = torch.tensor([(3 * i + 1) % R_rows for i in range(T)], dtype=torch.int32) if use_rows else None
    idx = rows.long() if rows is not None else torch.arange(T)
    summed = x0[idx].clone()
    for p in parts:
        summed += p[idx]
    want = R.rms_norm(summed, w.float(), 1e-5)
    wd = w.cuda()
    vp = C.c_void_p
    st = torch.cuda.current_stream().cuda_stream
    SENT16 = -32768
    for rk in range(world):
        owner = [int(idx[t]) % world for t in range(T)]
        rs = [torch.full((world, T_cap, H), -0.0, device="cuda") for _ in range(world)]      # every rank's slots
        outs = [torch.full((T + 1, H), 0, dtype=torch.int16, device="cuda").fill_(SENT16) for _ in range(world)]
        for t in range(T):
            if owner[t] == rk:
                for src in range(world):
                    if src != rk:
                        rs[rk][src, t] = (parts[src][idx[t]] + 0.0).cuda()   # "peers have pushed" (sanitised)
            else:
                outs[rk][t] = 0x3F80                                     # "the owner has broadcast" (bf16 1.0)
        x = x0.cuda()
        part = parts[rk].cuda()
        clear = torch.full((R_rows, H), 3.0, device="cuda")
        zero = torch.full((4096,), 9.0, device="cuda")
        reset = torch.zeros(2048, dtype=torch.int16, device="cuda")
        rc = lib.atx_allreduce_add_rmsnorm_ll(
            x.data_ptr(), part.data_ptr(), (vp * world)(*[r.data_ptr() for r in rs]),
            (vp * world)(*[o.data_ptr() for o in outs]), clear.data_ptr(), R_rows, reset.data_ptr(), reset.numel() * 2,
            T_cap, rk, world, rows.cuda().data_ptr() if rows is not None else None, wd.data_ptr(), T, H, 1e-5,
            zero.data_ptr(), zero.numel() * 4, 0, st)
        assert rc == 0, lib.atx_last_error().decode()
        torch.cuda.synchronize()
        assert torch.all(clear == 0) and torch.all(zero == 0) and torch.all(reset == SENT16)
        want_x = x0.clone()
        for t in range(T):
            o = owner[t]
            if o == rk:
                for p_ in range(world):     # produced here, delivered everywhere, bit-identical
                    got = outs[p_][t].view(torch.bfloat16).float().cpu()
                    assert torch.all((got - want[t]).abs() <= 2 ** -8 * want[t].abs() + 1e-6), (rk, t, p_)
                    assert torch.equal(outs[p_][t], outs[rk][t])
                assert torch.all(rs[rk][:, t].view(torch.int32) == -2 ** 31)      # slots re-armed
                if rows is None:
                    want_x[t] = summed[t]
            else:
                pushed = rs[o][rk, t].cpu()
                assert torch.equal(pushed, parts[rk][idx[t]] + 0.0)               # my partial, -0.0 sanitised
                for p_ in range(world):
                    if p_ != rk:
                        assert torch.all(outs[p_][t] == SENT16)                   # not mine to broadcast
        assert torch.all(outs[rk][T] == SENT16)
        assert torch.equal(x.cpu(), want_x)
    with pytest.raises(cabi.AtxError, match="co-resident"):
        cabi._check(lib.atx_allreduce_add_rmsnorm_ll(
            x.data_ptr(), part.data_ptr(), (vp * world)(*[r.data_ptr() for r in rs]),
            (vp * world)(*[o.data_ptr() for o in outs]), None, 0, None, 0, 4096, 0, world, None, wd.data_ptr(), 4096, H,
            1e-5, None, 0, 0, st))


def test_decode_attention_refuses_unsupported_shapes(cabi):
    """Pages of a multiple of 16 tokens, head_dim 64 | 128, at most 8 query heads per KV head: anything else is
    refused with the shape spelled out (there is no slower fallback kernel to fall into)."""
    kc, vc, bt, ks, vs = _paged_case(3, [40], 8, 2, 64)
    q = _bf16(torch.randn(1, 8 * 64)).cuda()
    out = torch.empty_like(q)
    sl = torch.tensor([40], dtype=torch.int32).cuda()
    with pytest.raises(cabi.AtxError, match="multiple of 16"):
        cabi.paged_attn_decode(q, kc.cuda(), vc.cuda(), bt.cuda(), sl, out, None, 1, 8, 2, 64, 8, 1, 0.125)
    with pytest.raises(cabi.AtxError, match="at most 8 query heads"):
        cabi.paged_attn_decode(q, kc.cuda(), vc.cuda(), bt.cuda(), sl, out, None, 1, 32, 2, 64, 16, 1, 0.125)


def test_prefill_attention_refuses_unsupported_shapes(cabi):
    """One prefill attention kernel (tcgen05, head_dim 64 | 128, 16-token pages): other shapes are refused, not
    routed to a slower path."""
    kc, vc, bt, ks, vs = _paged_case(4, [40], 8, 2, 64)
    q = _bf16(torch.randn(40, 8 * 64)).cuda()
    out = torch.empty_like(q)
    cu = torch.tensor([0, 40], dtype=torch.int32).cuda()
    ctx = torch.tensor([40], dtype=torch.int32).cuda()
    with pytest.raises(cabi.AtxError, match="unsupported head_dim"):
        cabi.paged_attn_prefill(q, kc.cuda(), vc.cuda(), bt.cuda(), cu, ctx, out, 1, 40, 8, 2, 64, 32, 0.125)
    with pytest.raises(cabi.AtxError, match="unsupported head_dim"):
        cabi.paged_attn_prefill(q, kc.cuda(), vc.cuda(), bt.cuda(), cu, ctx, out, 1, 40, 16, 4, 32, 16, 0.125)

