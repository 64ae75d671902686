"""Kernel micro-benchmarks on one B200 (CUDA events, L2-defeating weight rotation).

    python agentic-traffic-testing_b200/tools/microbench.py gemm [--T 6]
    python agentic-traffic-testing_b200/tools/microbench.py attn
Prints one JSON line per case: achieved GB/s (algorithmic bytes / event time).
"""
from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from atx import _cabi  # noqa: E402


def time_ms(fn, iters=20, warmup=3, graph=True):
    """Average device time of fn(). The `iters` calls are captured into ONE CUDA graph and replayed, so kernels
    shorter than a ctypes call (~10 us of host time) are timed at the device's pace, not the interpreter's."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if graph:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        g = torch.cuda.CUDAGraph()
        with torch.cuda.stream(side):
            with torch.cuda.graph(g, stream=side):
                for _ in range(iters):
                    fn()
            g.replay()   # warm the instantiated graph
            side.synchronize()
            s.record(side)
            g.replay()
            e.record(side)
            side.synchronize()
        return s.elapsed_time(e) / iters
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / iters


def bench_gemm(T: int, shapes=None, splits_list=(0,)):
    shapes = shapes or [("qkv", 6144, 4096), ("o", 4096, 4096), ("gate_up", 28672, 4096), ("down", 4096, 14336),
                        ("lm_head", 128256, 4096)]
    for name, N, K in shapes:
        copies = max(2, int(400e6 // (N * K * 2)) + 1)  # rotate through > 400 MB so L2 (126 MB) never helps
        ws = [torch.randn(N, K, device="cuda", dtype=torch.bfloat16) * 0.02 for _ in range(copies)]
        x = torch.randn(max(T, 16), K, device="cuda", dtype=torch.bfloat16)
        plans = [_cabi.GemmPlan(w, x) for w in ws]
        out = torch.zeros(max(T, 16), N, device="cuda")
        for splits in splits_list:
            mode = _cabi.GEMM_ADD_F32
            it = [0]

            def fn():
                plans[it[0] % copies].run(out, T, mode, splits)
                it[0] += 1

            ms = time_ms(fn, iters=4 * copies)
            byts = N * K * 2 + T * K * 2 + T * N * 4
            flops = 2.0 * T * N * K
            print(json.dumps({"kernel": "gemm", "name": name, "T": T, "N": N, "K": K,
                              "splits": splits or _cabi.gemm_auto_splits(N, K, T), "ms": round(ms, 4),
                              "GBps": round(byts / ms / 1e6, 1), "TFLOPs": round(flops / ms / 1e9, 1)}), flush=True)
        del plans, ws


def bench_attn():
    Hq, Hkv, D, bs = 32, 8, 128, 16
    for B, ctx, splits in [(6, 1152, 1), (6, 1152, 4), (6, 1152, 8), (16, 32768, 8), (16, 32768, 16), (64, 4096, 4)]:
        nblk = B * ((ctx + bs - 1) // bs)
        if nblk * Hkv * bs * D * 2 * 2 > 60e9:
            continue
        kc = torch.randn(nblk, Hkv, bs, D, device="cuda", dtype=torch.bfloat16)
        vc = torch.randn_like(kc)
        bt = torch.randperm(nblk, device="cuda").to(torch.int32).view(B, -1)
        sl = torch.full((B,), ctx, dtype=torch.int32, device="cuda")
        q = torch.randn(B, Hq * D, device="cuda", dtype=torch.bfloat16)
        out = torch.empty_like(q)
        ws = torch.empty(max(1, _cabi.decode_workspace_bytes(B, Hq, D, splits) // 4), device="cuda")
        ms = time_ms(lambda: _cabi.paged_attn_decode(q, kc, vc, bt, sl, out, ws, B, Hq, Hkv, D, bs, splits, D ** -0.5))
        byts = B * ctx * Hkv * D * 2 * 2
        print(json.dumps({"kernel": "attn_decode", "B": B, "ctx": ctx, "splits": splits, "ms": round(ms, 4),
                          "GBps": round(byts / ms / 1e6, 1)}), flush=True)


def bench_prefill_attn():
    Hq, Hkv, D, bs = 32, 8, 128, 16
    for S, L in [(6, 1024), (4, 8192), (2, 32768)]:
        nblk = S * (L // bs)
        kc = torch.randn(nblk, Hkv, bs, D, device="cuda", dtype=torch.bfloat16)
        vc = torch.randn_like(kc)
        bt = torch.arange(nblk, device="cuda", dtype=torch.int32).view(S, -1)
        q = torch.randn(S * L, Hq * D, device="cuda", dtype=torch.bfloat16)
        out = torch.empty_like(q)
        cu = torch.arange(0, (S + 1) * L, L, device="cuda", dtype=torch.int32)
        ctx = torch.full((S,), L, device="cuda", dtype=torch.int32)
        ms = time_ms(lambda: _cabi.paged_attn_prefill(q, kc, vc, bt, cu, ctx, out, S, L, Hq, Hkv, D, bs, D ** -0.5),
                     iters=5, warmup=2)
        flops = 4.0 * S * Hq * D * (L * (L + 1) / 2)   # causal
        print(json.dumps({"kernel": "attn_prefill", "S": S, "L": L, "ms": round(ms, 3),
                          "TFLOPs_causal": round(flops / ms / 1e9, 1)}), flush=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("what", choices=["gemm", "attn", "prefill", "all"])
    ap.add_argument("--T", type=int, nargs="*", default=[6])
    ap.add_argument("--splits", type=int, nargs="*", default=[0])
    ap.add_argument("--shapes", default="8b", choices=["8b", "70b_tp8"])
    a = ap.parse_args()
    if a.what in ("gemm", "all"):
        shapes = None if a.shapes == "8b" else [("qkv", 1280, 8192), ("o", 8192, 1024), ("gate_up", 7168, 8192),
                                                 ("down", 8192, 3584), ("lm_head", 16032, 8192)]
        for T in a.T:
            bench_gemm(T, shapes=shapes, splits_list=tuple(a.splits))
    if a.what in ("attn", "all"):
        bench_attn()
    if a.what in ("prefill", "all"):
        bench_prefill_attn()
