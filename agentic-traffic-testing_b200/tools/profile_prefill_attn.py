"""A few launches of the tcgen05 prefill attention at S=4 x 8192 tokens (8B head shapes) for ncu:
    ncu --set full --clock-control none --import-source on -k regex:attn_prefill_tc -s 2 -c 1 -o prof python <this file>
"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from atx import _cabi  # noqa: E402

_cabi.load()
Hq, Hkv, D, bs = 32, 8, 128, 16
S, L = 4, 8192
nblk = S * (L // bs)
kc = torch.randn(nblk, Hkv, bs, D, device="cuda", dtype=torch.bfloat16)
vc = torch.randn_like(kc)
bt = torch.arange(nblk, device="cuda", dtype=torch.int32).view(S, -1)
q = torch.randn(S * L, Hq * D, device="cuda", dtype=torch.bfloat16)
out = torch.empty_like(q)
cu = torch.arange(0, (S + 1) * L, L, device="cuda", dtype=torch.int32)
ctx = torch.full((S,), L, device="cuda", dtype=torch.int32)
for _ in range(3):
    _cabi.paged_attn_prefill(q, kc, vc, bt, cu, ctx, out, S, L, Hq, Hkv, D, bs, D ** -0.5)
torch.cuda.synchronize()
if len(sys.argv) > 1:      # any argument: CUDA-event timing instead of a run for ncu
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        _cabi.paged_attn_prefill(q, kc, vc, bt, cu, ctx, out, S, L, Hq, Hkv, D, bs, D ** -0.5)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    flops = 4 * D * Hq * S * L * (L + 1) / 2
    print(f"S={S} x {L}: {ms:.3f} ms = {flops / ms / 1e9:.0f} TFLOP/s causal")
