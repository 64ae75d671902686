"""A few launches of the persistent tcgen05 GEMM (gemm_wt_persistent_kernel<256>) at prefill sizes of the 8B model
for ncu: the gate||up projection with the SwiGLU epilogue and the down projection, T = 2304 (one cfg2 burst with
the prefix cache) and T = 8192 (one chunk of a long prompt). With any argument: CUDA-event timing instead.

    ncu --set full --clock-control none --import-source on -k regex:gemm_wt_persistent -c 4 -o prof python <this file>
"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from atx import _cabi  # noqa: E402

_cabi.load()
H, I = 4096, 14336
g = torch.Generator(device="cuda").manual_seed(0)
wgu = (torch.randn(2 * I, H, device="cuda", generator=g) * 0.02).to(torch.bfloat16)
wdown = (torch.randn(H, I, device="cuda", generator=g) * 0.02).to(torch.bfloat16)
timing = len(sys.argv) > 1
for T in (2304, 8192):
    x = torch.randn(T, H, device="cuda", generator=g).to(torch.bfloat16)
    act = torch.empty(T, I, device="cuda", dtype=torch.bfloat16)
    h = torch.zeros(T, H, device="cuda", dtype=torch.float32)
    p_gu, p_down = _cabi.GemmPlan(wgu, x), _cabi.GemmPlan(wdown, act)
    for rep in range(3 if timing else 1):
        p_gu.run(act, T, _cabi.GEMM_SWIGLU_BF16, 1)
        p_down.run(h, T, _cabi.GEMM_ADD_F32, 1)
    torch.cuda.synchronize()
    if timing:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            p_gu.run(act, T, _cabi.GEMM_SWIGLU_BF16, 1)
            p_down.run(h, T, _cabi.GEMM_ADD_F32, 1)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print(f"T={T}: gate||up + down {ms:.3f} ms = {2 * T * 3 * I * H / ms / 1e9:.0f} TFLOP/s")
