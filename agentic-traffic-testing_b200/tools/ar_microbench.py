"""Latency of the tensor-parallel all-reduce + residual add + RMSNorm step alone (no GEMMs between), per mode:

    sharded  atx_reduce_scatter_rmsnorm_all_gather  two-shot (reduce-scatter by row, all-gather by peer stores)
    nccl     dist.all_reduce + atx_rmsnorm(add=)    NCCL over NVLink + the local add/norm kernel
    ll       atx_allreduce_add_rmsnorm_ll           push protocol, the data is the flag (run last: it arms xn)

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 \
        agentic-traffic-testing_b200/tools/ar_microbench.py --T 32 --H 8192

Each mode is captured as one CUDA graph of --n calls (ping-pong buffers as the engine uses them) and replayed;
the time is device-timed, max over ranks. Rank 0 prints one JSON line.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from atx import _cabi  # noqa: E402


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--T", type=int, default=32)
    ap.add_argument("--H", type=int, default=8192)
    ap.add_argument("--n", type=int, default=160)
    ap.add_argument("--reps", type=int, default=20)
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    os.environ.setdefault("NCCL_DEBUG", "WARN")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    import torch.distributed._symmetric_memory as symm

    lib = _cabi.load()
    T, H = a.T, a.H
    part = symm.empty((2, T, H), dtype=torch.float32, device=dev)
    part.zero_()
    hp = symm.rendezvous(part, dist.group.WORLD)
    flags = symm.empty((64,), dtype=torch.int32, device=dev)
    flags.zero_()
    hf = symm.rendezvous(flags, dist.group.WORLD)
    xn = symm.empty((2, T, H), dtype=torch.bfloat16, device=dev)     # two parities for the low-latency form
    xn.view(torch.int16).fill_(-32768)
    hx = symm.rendezvous(xn, dist.group.WORLD)
    rs = symm.empty((2, world, T, H), dtype=torch.float32, device=dev)
    rs.fill_(-0.0)
    hr = symm.rendezvous(rs, dist.group.WORLD)
    vp = C.c_void_p
    parts = [(vp * world)(*[int(p) + b * T * H * 4 for p in hp.buffer_ptrs]) for b in (0, 1)]
    fl = (vp * world)(*[int(p) for p in hf.buffer_ptrs])
    outs = (vp * world)(*[int(p) for p in hx.buffer_ptrs])
    outs2 = [(vp * world)(*[int(p) + b * T * H * 2 for p in hx.buffer_ptrs]) for b in (0, 1)]
    rss = [(vp * world)(*[int(p) + b * world * T * H * 4 for p in hr.buffer_ptrs]) for b in (0, 1)]
    state = torch.zeros(2, dtype=torch.int32, device=dev)
    x = torch.randn(T, H, device=dev)
    w = torch.ones(H, dtype=torch.bfloat16, device=dev)
    local_part = torch.zeros(T, H, device=dev)
    torch.cuda.synchronize()
    dist.barrier()
    stream = torch.cuda.Stream(device=dev)

    def ck(rc):
        if rc != 0:
            raise RuntimeError(lib.atx_last_error().decode())

    def body(mode: str) -> None:
        st = torch.cuda.current_stream().cuda_stream
        for k in range(a.n):
            b = k % 2
            if mode == "sharded":
                ck(lib.atx_reduce_scatter_rmsnorm_all_gather(x.data_ptr(), parts[b], fl, outs, part[1 - b].data_ptr(), T,
                                                             state.data_ptr(), rank, world, None, w.data_ptr(), T, H,
                                                             1e-5, None, 0, _cabi.FLAG_PDL, st))
            elif mode == "ll":
                ck(lib.atx_allreduce_add_rmsnorm_ll(x.data_ptr(), part[b].data_ptr(), rss[b], outs2[b],
                                                    part[1 - b].data_ptr(), T, xn[1 - b].data_ptr(), T * H * 2, T, rank,
                                                    world, None, w.data_ptr(), T, H, 1e-5, None, 0, _cabi.FLAG_PDL, st))
            else:
                dist.all_reduce(local_part)
                ck(lib.atx_rmsnorm(x.data_ptr(), local_part.data_ptr(), None, w.data_ptr(), xn[0].data_ptr(), T, H, 1e-5,
                                   None, 0, _cabi.FLAG_PDL, st))

    res = {}
    for mode in ("sharded", "nccl", "ll"):
        with torch.cuda.stream(stream):
            if mode == "ll":   # the other modes left plain data in the activation buffers: arm them
                xn.view(torch.int16).fill_(-32768)
                stream.synchronize()
                dist.barrier()
            body(mode)   # warm-up outside capture
            stream.synchronize()
            dist.barrier()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=stream, capture_error_mode="thread_local"):
                body(mode)
            g.replay()
            stream.synchronize()
            dist.barrier()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(stream)
            for _ in range(a.reps):
                g.replay()
            e.record(stream)
            stream.synchronize()
            t = torch.tensor([s.elapsed_time(e)], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            res[mode] = round(float(t) * 1e3 / (a.reps * a.n), 2)
            del g
            x.normal_()   # the residual grows with every add: keep it finite
            torch.cuda.synchronize()
            dist.barrier()
    if rank == 0:
        print(json.dumps({"tool": "ar_microbench", "world": world, "T": T, "H": H, "calls_per_graph": a.n,
                          "us_per_call": res, "message_bytes_fp32": T * H * 4}), flush=True)
    torch.cuda.synchronize()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
