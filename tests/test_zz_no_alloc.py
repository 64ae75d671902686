"""The allocation-free decode-attention entry point (collected last on purpose: see the file name).

`atx_paged_attn_decode_static` serves callers that cannot allocate device memory — a process whose
HBM is already fully committed, one under a fixed memory reservation, a request path that must not
pay cudaMalloc's device synchronisation, or a binding from a runtime without a CUDA allocator. Its
split-KV scratch is an 8 MiB array inside the library, shared by all calls and never cleared, so the
checks are: same result as the caller-scratch entry point, on two different calls in a row (the
second must not be disturbed by what the first left behind), and a clear refusal beyond 8 MiB.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(seed, lens, Hq, Hkv, D, bs=16):
    g = torch.Generator().manual_seed(seed)
    need = [(L + bs - 1) // bs for L in lens]
    nblocks = sum(need) + 3
    perm = torch.randperm(nblocks, generator=g).tolist()
    kc = torch.randn(nblocks, Hkv, bs, D, generator=g).to(torch.bfloat16)
    vc = torch.randn(nblocks, Hkv, bs, D, generator=g).to(torch.bfloat16)
    width = max(need) + 1
    bt = torch.tensor([[perm.pop() for _ in range(n)] + [0] * (width - n) for n in need], dtype=torch.int32)
    q = torch.randn(len(lens), Hq * D, generator=g).to(torch.bfloat16)
    return q.cuda(), kc.cuda(), vc.cuda(), bt.cuda(), torch.tensor(lens, dtype=torch.int32).cuda()


def _both(cabi, case, Hq, Hkv, D, splits):
    q, kc, vc, bt, sl = case
    B = q.shape[0]
    ref = torch.empty_like(q)
    ws = torch.empty(max(1, cabi.decode_workspace_bytes(B, Hq, D, splits) // 4), device="cuda")
    cabi.paged_attn_decode(q, kc, vc, bt, sl, ref, ws, B, Hq, Hkv, D, 16, splits, D ** -0.5)
    got = torch.full_like(q, float("nan"))
    cabi.paged_attn_decode_static(q, kc, vc, bt, sl, got, B, Hq, Hkv, D, 16, splits, D ** -0.5)
    torch.cuda.synchronize()
    return got.float(), ref.float()


def test_static_scratch_decode_matches_caller_scratch_twice_in_a_row():
    from atx import _cabi as cabi

    cabi.load()
    assert cabi.decode_static_capacity() == 8 << 20
    # two calls back to back with different shapes, lengths and split counts: the second runs over
    # whatever the first left in the shared array
    first = _both(cabi, _case(101, [1, 17, 300, 1153, 64, 999], 32, 8, 128), 32, 8, 128, splits=8)
    second = _both(cabi, _case(202, [40, 5, 700], 8, 2, 64), 8, 2, 64, splits=3)
    third = _both(cabi, _case(303, [128, 129], 32, 8, 128), 32, 8, 128, splits=1)   # no scratch at all
    for got, ref in (first, second, third):
        assert torch.isfinite(got).all()
        # same tolerance as tests/test_gpu_kernels.py::test_paged_decode_attention
        assert torch.all((got - ref).abs() <= 2 ** -7 * ref.abs() + 2e-3), float((got - ref).abs().max())


def test_static_scratch_decode_matches_the_oracle_directly():
    """The same entry point against the CPU oracle itself (oracle/llama_ref.py:attention = modeling_llama.py:199-222),
    so that this file stands on its own: K/V of every sequence are read back through the block table."""
    from atx import _cabi as cabi
    from oracle import llama_ref as R

    cabi.load()
    lens, Hq, Hkv, D, bs = [1, 17, 300, 1153, 64, 999], 32, 8, 128, 16
    q, kc, vc, bt, sl = _case(404, lens, Hq, Hkv, D)
    got = torch.full_like(q, float("nan"))
    cabi.paged_attn_decode_static(q, kc, vc, bt, sl, got, len(lens), Hq, Hkv, D, bs, 8, D ** -0.5)
    torch.cuda.synchronize()
    got = got.float().cpu().view(len(lens), Hq, D)
    for b, L in enumerate(lens):
        blocks = bt[b, :(L + bs - 1) // bs].long()
        k = kc[blocks].permute(0, 2, 1, 3).reshape(-1, Hkv, D)[:L].float().cpu()
        v = vc[blocks].permute(0, 2, 1, 3).reshape(-1, Hkv, D)[:L].float().cpu()
        want = R.attention(q[b].float().cpu().view(1, Hq, D), k, v, L - 1)[0]
        # same tolerance as tests/test_gpu_kernels.py::test_paged_decode_attention
        assert torch.all((got[b] - want).abs() <= 2 ** -7 * want.abs() + 2e-3), (b, L)


def test_static_scratch_decode_refuses_oversize_requests_clearly():
    from atx import _cabi as cabi

    cabi.load()
    q, kc, vc, bt, sl = _case(7, [32] * 64, 32, 8, 128)          # 64 x 32 heads x 32 splits x 130 floats = 34 MB
    out = torch.empty_like(q)
    with pytest.raises(cabi.AtxError) as e:
        cabi.paged_attn_decode_static(q, kc, vc, bt, sl, out, 64, 32, 8, 128, 16, 32, 128 ** -0.5)
    msg = str(e.value)
    assert "atx error -4" in msg and "built-in array holds 8388608" in msg and "num_splits=32" in msg
    # a request that fits still works afterwards
    cabi.paged_attn_decode_static(q, kc, vc, bt, sl, out, 64, 32, 8, 128, 16, 2, 128 ** -0.5)
    torch.cuda.synchronize()
    assert torch.isfinite(out.float()).all()
