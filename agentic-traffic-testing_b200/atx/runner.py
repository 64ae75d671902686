"""ModelRunner: executes one scheduler step (a flat batch of decode tokens + prompt chunks) on one
B200 by calling the C-ABI kernels; owns the weights, the paged KV pools and every activation buffer.

This is the part of vLLM's model runner that the reference reaches through
`AsyncLLMEngine.generate` (llm/serve_llm.py:527), rebuilt B200-first:
  * all buffers are allocated once (no allocator traffic on the hot path, CUDA-graph safe),
  * decode-only steps replay a CUDA graph per (batch bucket, KV-split bucket),
  * every kernel is launched with programmatic dependent launch so that a GEMM's weight
    prefetch overlaps the tail of the kernel before it,
  * host inputs go through one pinned staging buffer; the only device->host traffic per step is
    the sampled token ids.
Layer schedule (per token batch of T rows; h is the fp32 residual stream):
  xn = rmsnorm(h)            -> bf16      (also clears the split-K accumulator that follows)
  qkv = xn @ Wqkv^T          -> fp32      (tcgen05 GEMM, split-K for decode)
  q, K/V pages = rope(qkv)   -> bf16      (scatter by slot mapping)
  a = paged attention        -> bf16
  h += a @ Wo^T                            (GEMM epilogue adds into the residual)
  xn = rmsnorm(h); act = swiglu(xn @ Wgu^T) (GEMM epilogue, bf16); h += act @ Wdown^T
Tensor parallel (tp_size > 1): the two `h +=` projections produce partial sums, and the all-reduce, the
residual add and the following rmsnorm are ONE kernel of ours over NVLink peer memory (`_fused_norm`:
low-latency push form for decode-sized steps, flag-barrier form up to 256 rows, NCCL beyond).
"""
from __future__ import annotations

import ctypes as C
import math
import os
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from . import _cabi
from .config import EngineConfig, ModelConfig
from .scheduler import Chunk, StepPlan
from .weights import ModelWeights, rope_table

_GRAPH_BUCKETS = [1, 2, 4, 6, 8, 12, 16, 24, 32, 48, 64]


def _bucket(n: int, buckets: List[int]) -> Optional[int]:
    for b in buckets:
        if n <= b:
            return b
    return None


class ModelRunner:
    def __init__(self, mcfg: ModelConfig, ecfg: EngineConfig, weights: ModelWeights):
        if not torch.cuda.is_available():
            raise RuntimeError("atx ModelRunner needs a CUDA device (there is no CPU fallback)")
        self.lib = _cabi.load()
        self.m, self.e, self.w = mcfg, ecfg, weights
        self.dev = weights.device
        torch.cuda.set_device(self.dev)
        m, e = mcfg, ecfg
        self.Tmax = max(16, e.max_num_batched_tokens)
        self.Smax = max(1, e.max_num_seqs)
        self.bs = e.block_size
        if self.bs != 16:
            raise ValueError(f"block_size={self.bs}: KV pages hold 16 tokens (vLLM's default; the prefill attention "
                             "loads one page of one KV head per TMA box)")
        self.MB = (e.max_model_len + self.bs - 1) // self.bs
        self.tp, self.tp_rank, self.tp_group = max(1, e.tp_size), e.tp_rank, None
        tp = self.tp
        if tp > 1:
            import torch.distributed as dist

            if not dist.is_initialized() or dist.get_world_size() % tp:
                raise RuntimeError("tensor parallelism needs an initialised NCCL process group (torchrun)")
            if m.num_key_value_heads % tp or m.intermediate_size % (64 * tp) or m.vocab_size % tp:
                raise ValueError(f"tp_size={tp} must divide kv heads, intermediate_size/64 and the vocabulary")
            first = dist.get_rank() // tp * tp
            self.tp_group = dist.new_group(list(range(first, first + tp))) if dist.get_world_size() != tp else None
        H, I, V = m.hidden_size, m.intermediate_size // tp, m.vocab_size
        q_size = m.q_size // tp
        qkv_size = m.qkv_size // tp
        bf, f32, i32 = torch.bfloat16, torch.float32, torch.int32
        dev = self.dev
        T, S = self.Tmax, max(self.Smax, 16)
        self.h = torch.zeros(T, H, dtype=f32, device=dev)
        self.qkv = torch.zeros(T, qkv_size, dtype=f32, device=dev)
        self.q = torch.zeros(T, q_size, dtype=bf, device=dev)
        self.attn = torch.zeros(T, q_size, dtype=bf, device=dev)
        self.act = torch.zeros(T, I, dtype=bf, device=dev)
        self.part = torch.zeros(T, H, dtype=f32, device=dev) if tp > 1 else None  # partial o/down sums (TP)
        # ... of prefill-sized steps: stored and all-reduced in bf16 (half the NCCL bytes; what vLLM reduces too)
        self.part_bf = torch.zeros(T, H, dtype=bf, device=dev) if tp > 1 else None
        self.allreduce_calls = 0
        self.fused_ar = False      # all-reduce + add + norm in our own kernels over peer memory (else: NCCL + norm kernel)
        self.xn = self.xn2 = self.xl = None
        if e.tp_allreduce not in ("auto", "nccl"):
            raise ValueError(f"tp_allreduce={e.tp_allreduce!r}: expected auto | nccl")
        if tp > 1 and e.tp_allreduce != "nccl":
            self._setup_fused_allreduce(T, H, S)
        if self.xn is None:   # (the two-shot kernels need them in peer-mapped memory: allocated above)
            self.xn = torch.zeros(T, H, dtype=bf, device=dev)
            self.xl = torch.zeros(S, H, dtype=bf, device=dev)
        if self.xn2 is None:
            self.xn2 = self.xn    # input of gate_up; a buffer of its own only under the low-latency all-reduce
        self.logits = torch.zeros(S, V, dtype=f32, device=dev)
        self.out_ids = torch.zeros(S, dtype=i32, device=dev)
        if tp > 1:   # (score, global index) of this rank's vocabulary shard per row, and every rank's after the exchange
            self.pairs = torch.zeros(S, 2, dtype=f32, device=dev)
            self.pairs_all = torch.zeros(tp, S, 2, dtype=f32, device=dev)
        self.out_host = [torch.zeros(S, dtype=i32).pin_memory() for _ in range(2)]  # ping-pong: one step may be in flight
        self._out_events = [torch.cuda.Event(), torch.cuda.Event()]
        self._out_flip = 0
        self.rope = rope_table(m, e.max_model_len, dev)
        self.max_splits = 32
        self.ws = torch.zeros(max(1, _cabi.decode_workspace_bytes(S, m.num_attention_heads // tp, m.head_dim,
                                                                  self.max_splits) // 4), dtype=f32, device=dev)

        # ---- staged integer / float inputs: one pinned host mirror, one device buffer -----------
        S_ = self.Smax
        off, self._off = 0, {}

        def region(name: str, n: int) -> None:
            nonlocal off
            off = (off + 3) // 4 * 4  # 16-byte aligned regions
            self._off[name] = (off, n)
            off += n

        for name, n in (("ids", T), ("pos", T), ("slots", T), ("dec_lens", S_), ("pf_cu", S_ + 1), ("pf_ctx", S_),
                        ("rows", S_), ("steps", S_), ("temps", S_), ("top_p", S_), ("seeds", 2 * S_),
                        ("dec_bt", S_ * self.MB), ("pf_bt", S_ * self.MB)):
            region(name, n)
        self._meta_lo = self._off["dec_lens"][0]
        self._meta_hi = self._off["seeds"][0] + self._off["seeds"][1]
        self.hbuf = torch.zeros(off, dtype=i32).pin_memory()
        self.dbuf = torch.zeros(off, dtype=i32, device=dev)
        self.hnp = self.hbuf.numpy()
        self._hv = {k: self.hnp[o:o + n] for k, (o, n) in self._off.items()}
        self._hv["temps_f"] = self._hv["temps"].view(np.float32)
        self._hv["top_p_f"] = self._hv["top_p"].view(np.float32)
        self._hv["seeds_q"] = self._hv["seeds"].view(np.int64)
        self._hv["dec_bt2"] = self._hv["dec_bt"].reshape(S_, self.MB)
        self._hv["pf_bt2"] = self._hv["pf_bt"].reshape(S_, self.MB)
        base = self.dbuf.data_ptr()
        self._dp = {k: base + 4 * o for k, (o, n) in self._off.items()}

        # ---- GEMM plans (TMA descriptors) --------------------------------------------------------
        self.plans = []
        for l in weights.layers:
            self.plans.append((_cabi.GemmPlan(l.wqkv, self.xn), _cabi.GemmPlan(l.wo, self.attn),
                               _cabi.GemmPlan(l.wgu, self.xn2), _cabi.GemmPlan(l.wdown, self.act)))
        self.plan_lm = _cabi.GemmPlan(weights.lm_head, self.xl)

        # ---- paged KV pools, sized from what is left of the HBM budget ---------------------------
        self.block_bytes = 2 * m.num_hidden_layers * (m.num_key_value_heads // tp) * self.bs * m.head_dim * 2
        if e.num_gpu_blocks is not None:
            nb = int(e.num_gpu_blocks)
        else:
            free, total = torch.cuda.mem_get_info(dev)
            budget = int(total * e.gpu_memory_utilization) - (total - free) - (512 << 20)
            nb = budget // self.block_bytes
            if nb < self.MB:
                raise RuntimeError(f"KV budget {budget / 2**30:.1f} GiB holds {nb} blocks; one sequence of "
                                   f"max_model_len={e.max_model_len} needs {self.MB}")
        self.num_blocks = nb
        L, Hkv, D = m.num_hidden_layers, m.num_key_value_heads // tp, m.head_dim
        if tp > 1 and e.num_gpu_blocks is None:
            # every rank must own the same number of blocks: the schedulers run in lockstep
            import torch.distributed as dist

            t_nb = torch.tensor([nb], device=dev)
            dist.all_reduce(t_nb, op=dist.ReduceOp.MIN, group=self.tp_group)
            nb = self.num_blocks = int(t_nb.item())
        self.k_cache = torch.zeros(L, nb, Hkv, self.bs, D, dtype=bf, device=dev)
        self.v_cache = torch.zeros(L, nb, Hkv, self.bs, D, dtype=bf, device=dev)

        self._graphs: Dict[Tuple[int, int, bool], torch.cuda.CUDAGraph] = {}
        self._graph_launches: Dict[Tuple[int, int, bool], int] = {}
        self.stream = torch.cuda.Stream(device=dev)   # every launch and copy of this runner
        self._h2d_done = torch.cuda.Event()
        self.flags = (_cabi.FLAG_PDL if e.use_pdl else 0) | (_cabi.FLAG_ORDERED if e.deterministic else 0)
        self.scale = float(m.head_dim) ** -0.5
        self.launches_last_step = 0
        self.kernel_launches = 0  # cumulative count of our kernels launched (graph replays included)
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        # bench hook: when a list, every step appends (start_event, end_event, n_decode, prefill_tokens, ctx_sum)
        self.step_events: Optional[list] = None

    # ----------------------------------------------------------------------------------------------
    def _ck(self, rc: int) -> None:
        if rc != 0:
            raise _cabi.AtxError(f"atx error {rc}: {self.lib.atx_last_error().decode()}")

    def _forward(self, T: int, n_dec: int, dec_splits: int, n_pf: int, max_q: int, n_rows: int,
                 whole_rows: bool = False) -> int:
        """Issue every kernel of one step on the current stream. Returns the launch count.

        Tensor parallel (tp > 1, Megatron split: q/k/v/gate/up by output features = heads, o/down by
        input features): o_proj and down_proj produce PARTIAL sums. For decode-sized steps the
        all-reduce, the residual add and the following RMSNorm are one kernel over NVLink peer memory
        (atx_allreduce_add_rmsnorm_ll up to 64 rows, atx_reduce_scatter_rmsnorm_all_gather up to 256; ping-pong
        symmetric buffers); prefill-sized steps store the partial in bf16, all-reduce it with NCCL's ring and fold it
        into the residual inside the RMSNorm kernel (`add`). The LM head is
        vocabulary-parallel: each rank computes its column slice of the logits, reduces every row of it to one
        (score, global index) pair (atx_sample_shard: arg-max, or Gumbel-max keyed by the global index), the ranks
        exchange the pairs (8 bytes per row and rank, one NCCL all-gather) and pick the best (atx_sample_merge):
        the same token on every rank, and the token the unsharded sampler draws. Only when a request asks for a
        nucleus (top_p < 1: `whole_rows`) are the rows themselves assembled (all-reduce of a zero-filled buffer)."""
        lib, m, w, ck = self.lib, self.m, self.w, self._ck
        st = torch.cuda.current_stream().cuda_stream
        fl = self.flags
        tp = self.tp
        H, V = m.hidden_size, m.vocab_size
        D = m.head_dim
        Hq, Hkv, I = m.num_attention_heads // tp, m.num_key_value_heads // tp, m.intermediate_size // tp
        q_size, qkv_size = Hq * D, (Hq + 2 * Hkv) * D
        dp = self._dp
        h, xn, qkv, q, attn, act = (t.data_ptr() for t in (self.h, self.xn, self.qkv, self.q, self.attn, self.act))
        fused = tp > 1 and self.fused_ar and T <= self.FUSED_AR_MAX_T
        self._ll_now = fused and T <= self.LL_AR_MAX_T
        xn2 = self.xn2.data_ptr()
        if self._ll_now:
            # steps that took another path may have left plain data in the first activation buffer: re-arm it
            # (the other one and xl are re-armed by the first all-reduce call below)
            self._arm_a.fill_(-32768)
        bf_part = False
        if tp == 1:
            targets = (h, h)                      # projections add straight into the residual stream
        elif fused:
            targets = (self.sym_part[0].data_ptr(), self.sym_part[1].data_ptr())
        else:
            # NCCL path. Decode-sized steps accumulate split-K partials in fp32; beyond 64 rows the GEMM is the
            # persistent kernel (no split-K), which stores its tile once: bf16, then a bf16 all-reduce
            bf_part = T > 64
            if not bf_part:
                self.part[:T].zero_()
            targets = (self.part_bf.data_ptr() if bf_part else self.part.data_ptr(),) * 2
        n = 0
        k = 0            # index of the next projection whose partial sum needs reducing (TP)
        pending = None   # None | buffer index (fused) | True (NCCL-reduced partial waiting in self.part)

        def norm(w_ptr: int, out_ptr: int, zero_ptr=None, zero_bytes: int = 0, rows_ptr=None, rows_n: int = T) -> None:
            nonlocal n
            if tp > 1 and fused and pending is not None:
                self._fused_norm(pending, T, w_ptr, out_ptr, rows_ptr, rows_n, zero_ptr, zero_bytes, st)
            else:
                add, afl = None, fl
                if tp > 1 and pending is not None:
                    add = self.part_bf.data_ptr() if bf_part else self.part.data_ptr()
                    afl = fl | (_cabi.FLAG_ADD_BF16 if bf_part else 0)
                ck(lib.atx_rmsnorm(h, add, rows_ptr, w_ptr, out_ptr, rows_n, H, m.rms_norm_eps, zero_ptr, zero_bytes, afl,
                                   st))
            n += 1

        def project(plan) -> None:
            """o_proj / down_proj: accumulate into the residual (tp 1) or into this round's partial buffer."""
            nonlocal n, k, pending
            if bf_part:
                ck(lib.atx_gemm_run(plan._h, targets[0], H, T, _cabi.GEMM_STORE_BF16, 1, fl, st))
            else:
                ck(lib.atx_gemm_run(plan._h, targets[k % 2], H, T, _cabi.GEMM_ADD_F32, 0, fl, st))
            n += 1
            if tp > 1:
                if fused:
                    pending = k % 2
                else:
                    self._all_reduce(self.part_bf[:T] if bf_part else self.part[:T])
                    pending = True
                k += 1

        ck(lib.atx_embed_gather(dp["ids"], w.embed.data_ptr(), h, T, H, V, st))
        n += 1
        sp_qkv = lib.atx_gemm_auto_splits(qkv_size, H, T)
        qb = q_size * 2  # bytes per row of q / attn
        for li, l in enumerate(w.layers):
            p_qkv, p_o, p_gu, p_down = self.plans[li]
            kc, vc = self.k_cache[li].data_ptr(), self.v_cache[li].data_ptr()
            if sp_qkv > 1:   # split-K accumulates into qkv: the norm kernel's spare CTAs clear it
                norm(l.ln1.data_ptr(), xn, qkv, T * qkv_size * 4)
                ck(lib.atx_gemm_run(p_qkv._h, qkv, qkv_size, T, _cabi.GEMM_ADD_F32, sp_qkv, fl, st))
            else:
                norm(l.ln1.data_ptr(), xn)
                ck(lib.atx_gemm_run(p_qkv._h, qkv, qkv_size, T, _cabi.GEMM_STORE_F32, 1, fl, st))
            n += 1
            ck(lib.atx_rope_kv_write(qkv, dp["pos"], dp["slots"], self.rope.data_ptr(), q, kc, vc, T, Hq, Hkv, D,
                                     self.bs, fl, st))
            n += 1
            if n_dec:
                ck(lib.atx_paged_attn_decode(q, kc, vc, dp["dec_bt"], dp["dec_lens"], attn, self.ws.data_ptr(),
                                             self.ws.numel() * 4, n_dec, Hq, Hkv, D, self.bs, self.MB, dec_splits,
                                             self.scale, fl, st))
                n += 2 if dec_splits > 1 else 1
            if n_pf:
                ck(lib.atx_paged_attn_prefill(q + n_dec * qb, kc, vc, dp["pf_bt"], dp["pf_cu"], dp["pf_ctx"],
                                              attn + n_dec * qb, n_pf, max_q, T - n_dec, self.num_blocks, Hq, Hkv, D,
                                              self.bs, self.MB, self.scale, fl, st))
                n += 1
            project(p_o)
            norm(l.ln2.data_ptr(), xn2)
            ck(lib.atx_gemm_run(p_gu._h, act, I, T, _cabi.GEMM_SWIGLU_BF16, 0, fl, st))  # silu(g)*u in the epilogue
            n += 1
            project(p_down)
        if n_rows:
            norm(w.norm.data_ptr(), self.xl.data_ptr(), rows_ptr=dp["rows"], rows_n=n_rows)
            if tp > 1 and not whole_rows:
                import torch.distributed as dist

                Vl = V // tp
                shard = self.logits.data_ptr() + 4 * Vl * self.tp_rank       # this rank's columns of the [S, V] buffer
                ck(lib.atx_gemm_run(self.plan_lm._h, shard, V, n_rows, _cabi.GEMM_STORE_F32, 1, fl, st))
                ck(lib.atx_sample_shard(shard, V, Vl * self.tp_rank, dp["temps"], dp["seeds"], dp["steps"],
                                        self.pairs.data_ptr(), n_rows, Vl, fl, st))
                dist.all_gather_into_tensor(self.pairs_all, self.pairs, group=self.tp_group)
                self.allreduce_calls += 1
                ck(lib.atx_sample_merge(self.pairs_all.data_ptr(), tp, self.pairs.shape[0], self.out_ids.data_ptr(), fl,
                                        st))
                return n + 3
            if tp > 1:
                Vl = V // tp
                self.logits[:n_rows].zero_()
                ck(lib.atx_gemm_run(self.plan_lm._h, self.logits.data_ptr() + 4 * Vl * self.tp_rank, V, n_rows,
                                    _cabi.GEMM_STORE_F32, 1, fl, st))
                self._all_reduce(self.logits[:n_rows])
            else:
                ck(lib.atx_gemm_run(self.plan_lm._h, self.logits.data_ptr(), V, n_rows, _cabi.GEMM_STORE_F32, 1, fl,
                                    st))
            ck(lib.atx_sample(self.logits.data_ptr(), V, dp["temps"], dp["top_p"], dp["seeds"], dp["steps"],
                              self.out_ids.data_ptr(), n_rows, V, fl, st))
            n += 2
        elif fused:
            # nothing samples this step, but the ping-pong invariant (one clean buffer) still has to advance
            norm(w.norm.data_ptr(), xn)
        return n

    # rows up to which a step uses the fused all-reduce kernels (beyond: NCCL's bandwidth-optimal ring)
    FUSED_AR_MAX_T = 256
    # rows up to which the low-latency push form is used: its whole grid spins on remote data and must be
    # co-resident (rows x H/2048 CTAs <= 2 x SMs)
    LL_AR_MAX_T = 64

    def _setup_fused_allreduce(self, T: int, H: int, S: int) -> None:
        """Symmetric (peer-mapped) memory for the two all-reduce + add + norm kernels: ping-pong partial buffers and
        flag arrays (flag-barrier form), the activation buffers xn2 / xl / xn (every rank writes its rows into every
        rank's copy), and the receive slots of the push form, armed with the sentinel. torch's symmetric-memory
        allocator is used only to obtain peer-mapped memory; the kernels are ours."""
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm

        group = self.tp_group if self.tp_group is not None else dist.group.WORLD
        rows = min(T, self.FUSED_AR_MAX_T)
        try:
            self.sym_part = symm.empty((2, rows, H), dtype=torch.float32, device=self.dev)
            self.sym_part.zero_()
            hp = symm.rendezvous(self.sym_part, group)
            self.sym_flags = symm.empty((64,), dtype=torch.int32, device=self.dev)
            self.sym_flags.zero_()
            hf = symm.rendezvous(self.sym_flags, group)
            # activation buffers every rank writes into: A = xn2 [T, H] (gate_up input), xl [S, H] (LM-head
            # input), B = xn [T, H] (qkv input); xl and the head of B are contiguous (one re-arm region)
            self.sym_x = symm.empty((2 * T + S, H), dtype=torch.bfloat16, device=self.dev)
            hx = symm.rendezvous(self.sym_x, group)
            cap = min(T, self.LL_AR_MAX_T)
            self.sym_rs = symm.empty((2, self.tp, cap, H), dtype=torch.float32, device=self.dev)
            self.sym_rs.fill_(-0.0)                                    # the sentinel: "nothing has arrived"
            hr = symm.rendezvous(self.sym_rs, group)
            self.sym_x.view(torch.int16).fill_(-32768)                 # bf16 -0.0
        except Exception as exc:  # no peer mapping on this box: keep the NCCL path, loudly
            print(f"[atx] symmetric memory unavailable ({exc!r}); tensor-parallel all-reduce stays on NCCL", flush=True)
            return
        stride = rows * H * 4
        vp = C.c_void_p
        self._ar_parts = [(vp * self.tp)(*[int(p) + b * stride for p in hp.buffer_ptrs]) for b in (0, 1)]
        self._ar_flags = (vp * self.tp)(*[int(p) for p in hf.buffer_ptrs])
        self._ar_state = torch.zeros(2, dtype=torch.int32, device=self.dev)
        self._ar_handles = (hp, hf, hx, hr)
        self.xn2, self.xl, self.xn = self.sym_x[:T], self.sym_x[T:T + S], self.sym_x[T + S:]
        self._ar_x_base = [int(p) for p in hx.buffer_ptrs]
        self._ar_out_ptrs = {}     # byte offset into sym_x -> array of the peers' pointers
        self._ll_cap = cap
        self._ll_rs = [(vp * self.tp)(*[int(p) + b * self.tp * cap * H * 4 for p in hr.buffer_ptrs]) for b in (0, 1)]
        # parity 0 writes A and re-arms xl + the head of B; parity 1 writes B (or xl) and re-arms the head of A
        self._ll_reset = [(self.xl.data_ptr(), (S + cap) * H * 2), (self.xn2.data_ptr(), cap * H * 2)]
        self._arm_a = self.sym_x[:cap].view(torch.int16)
        torch.cuda.synchronize(self.dev)
        dist.barrier(group=group)
        self.fused_ar = True

    def _fused_norm(self, buf: int, T: int, w_ptr: int, out_ptr: int, rows_ptr, n_rows: int, zero_ptr, zero_bytes: int,
                    st: int) -> None:
        """all-reduce(part[buf]) + residual add + RMSNorm; hands the other buffer back zeroed."""
        m = self.m
        other = self.sym_part[1 - buf].data_ptr()
        off = out_ptr - self.sym_x.data_ptr()
        outs = self._ar_out_ptrs.get(off)
        if outs is None:
            assert 0 <= off < self.sym_x.numel() * 2, "all-reduce output must live in the symmetric buffer"
            outs = self._ar_out_ptrs[off] = (C.c_void_p * self.tp)(*[b + off for b in self._ar_x_base])
        if self._ll_now:      # decode-sized: the push form
            in_a = off < self.xl.data_ptr() - self.sym_x.data_ptr()
            assert in_a == (buf == 0), "activation buffer does not match the all-reduce parity"
            reset_ptr, reset_bytes = self._ll_reset[buf]
            self._ck(self.lib.atx_allreduce_add_rmsnorm_ll(
                self.h.data_ptr(), self.sym_part[buf].data_ptr(), self._ll_rs[buf], outs, other, T, reset_ptr,
                reset_bytes, self._ll_cap, self.tp_rank, self.tp, rows_ptr, w_ptr, n_rows, m.hidden_size,
                m.rms_norm_eps, zero_ptr, zero_bytes, self.flags, st))
        else:                 # up to FUSED_AR_MAX_T rows: the flag-barrier form
            self._ck(self.lib.atx_reduce_scatter_rmsnorm_all_gather(
                self.h.data_ptr(), self._ar_parts[buf], self._ar_flags, outs, other, T, self._ar_state.data_ptr(),
                self.tp_rank, self.tp, rows_ptr, w_ptr, n_rows, m.hidden_size, m.rms_norm_eps, zero_ptr, zero_bytes,
                self.flags, st))
        self.allreduce_calls += 1

    def allreduce_alone_us(self, T: int, n_calls: int, reps: int = 10) -> Optional[float]:
        """Bench hook (BASELINE configs[3] asks for the all-reduce share of a step): the all-reduce + residual add +
        RMSNorm calls of one T-row step back to back with nothing between them, captured as one CUDA graph and
        replayed; microseconds per call on this rank. Every rank of the group must call it. The residual stream is
        left meaningless (call it between requests only)."""
        if not (self.tp > 1 and self.fused_ar and T <= self.FUSED_AR_MAX_T) or n_calls % 2:
            return None
        import torch.distributed as dist

        ln = self.w.layers[0].ln1.data_ptr()

        def body() -> None:
            st = torch.cuda.current_stream().cuda_stream
            self._ll_now = T <= self.LL_AR_MAX_T
            if self._ll_now:
                self._arm_a.fill_(-32768)
            for k in range(n_calls):      # the parities alternate exactly as in a step: o_proj -> xn2, down_proj -> xn
                self._fused_norm(k % 2, T, ln, (self.xn2 if k % 2 == 0 else self.xn).data_ptr(), None, T, None, 0, st)

        s = torch.cuda.Stream(device=self.dev)
        torch.cuda.synchronize(self.dev)
        dist.barrier(group=self.tp_group)
        with torch.cuda.stream(s):
            body()
            s.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=s, capture_error_mode="thread_local"):
                body()
            g.replay()
            s.synchronize()
            dist.barrier(group=self.tp_group)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                g.replay()
            e1.record()
            s.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / (reps * n_calls)
        del g
        torch.cuda.synchronize(self.dev)
        return us

    def _all_reduce(self, t: torch.Tensor) -> None:
        import torch.distributed as dist

        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.tp_group)
        self.allreduce_calls += 1

    # ----------------------------------------------------------------------------------------------
    def _h2d(self, name: str, count: int) -> None:
        o, _ = self._off[name]
        self.dbuf[o:o + count].copy_(self.hbuf[o:o + count], non_blocking=True)
        self.h2d_bytes += 4 * count

    def decode_splits(self, n_dec: int, max_len: int) -> int:
        """KV partitions per sequence. Short-context decode attention is a latency chain (block table
        -> K/V rows -> softmax), so cut it short: ~64 tokens per CTA while the grid stays within a
        few waves (<= 16 CTAs of 128 threads per SM), power-of-two buckets for the graph cache."""
        if n_dec == 0:
            return 1
        by_len = max(1, (max_len + 63) // 64)
        by_grid = max(1, (16 * 148) // max(1, n_dec * max(1, self.m.num_key_value_heads // self.tp)))
        s = 1
        while s * 2 <= min(by_len, by_grid, self.max_splits):
            s *= 2
        return s

    def execute(self, plan: StepPlan) -> Dict[int, int]:
        """Run one step and wait for it; returns {Sequence.uid: sampled token id} for every chunk that samples."""
        h = self.launch(plan)
        return self.wait(h) if h is not None else {}

    def launch(self, plan: StepPlan, chained: bool = False):
        """Enqueue one step (staging copies, kernels or graph replay, token read-back) without waiting.
        `chained`: the step directly follows an in-flight decode step over the SAME sequences in the same
        order; its input token ids are not known to the host yet and are taken on the device from the
        previous step's sampler output (the host fills everything else: positions, slots, lengths)."""
        with torch.cuda.stream(self.stream):
            return self._launch(plan, chained)

    def wait(self, handle) -> Dict[int, int]:
        """Block until the step's sampled ids have reached the pinned host buffer."""
        if handle is None:
            return {}
        sampling, n_dec, B, buf, ev = handle
        if ev is None:
            return {}
        ev.synchronize()
        oh = self.out_host[buf].numpy()
        out: Dict[int, int] = {}
        for r, c in enumerate(sampling):
            slot_r = r if r < n_dec else (B - n_dec) + r
            out[c.seq.uid] = int(oh[slot_r])
        return out

    def _launch(self, plan: StepPlan, chained: bool):
        self._h2d_done.synchronize()  # the previous step's staging copies have left the pinned buffer
        hv, bs = self._hv, self.bs
        decs, pfs = plan.decodes, plan.prefills
        n_dec, n_pf = len(decs), len(pfs)
        T_real = n_dec + sum(c.count for c in pfs)
        if T_real == 0:
            return None
        if T_real > self.Tmax or n_dec + n_pf > self.Smax:
            raise RuntimeError("step exceeds the runner's static capacity")
        # TP steps contain NCCL all-reduces: they capture fine (thread-local capture mode), but the graphs
        # must be dropped before the process group is destroyed (close()).
        use_graph = self.e.enable_cuda_graphs and n_pf == 0 and _bucket(n_dec, _GRAPH_BUCKETS) is not None \
            and _bucket(n_dec, _GRAPH_BUCKETS) <= self.Smax
        B = _bucket(n_dec, _GRAPH_BUCKETS) if use_graph else n_dec
        T = B + (T_real - n_dec)

        ids, pos, slots = hv["ids"], hv["pos"], hv["slots"]
        rows: List[int] = []
        sampling: List[Chunk] = []
        max_len = 0
        for i, c in enumerate(decs):
            s = c.seq
            p = c.start
            ids[i] = 0 if chained else s.tokens[p]   # chained: filled on the device from the previous step
            pos[i] = p
            tbl = s.blocks.table
            slots[i] = tbl[p // bs] * bs + p % bs
            hv["dec_bt2"][i, :len(tbl)] = tbl
            hv["dec_lens"][i] = p + 1
            max_len = max(max_len, p + 1)
            rows.append(i)
            sampling.append(c)
        for i in range(n_dec, B):  # graph padding: no KV write, empty attention, ignored sample
            ids[i] = 0
            pos[i] = 0
            slots[i] = -1
            hv["dec_lens"][i] = 0
            rows.append(i)
        o = B
        max_q = 0
        hv["pf_cu"][0] = 0
        for j, c in enumerate(pfs):
            s = c.seq
            a, n = c.start, c.count
            ids[o:o + n] = s.tokens[a:a + n]
            pr = np.arange(a, a + n, dtype=np.int32)
            pos[o:o + n] = pr
            tbl = np.asarray(s.blocks.table, dtype=np.int32)
            slots[o:o + n] = tbl[pr // bs] * bs + pr % bs
            hv["pf_bt2"][j, :len(tbl)] = tbl
            hv["pf_cu"][j + 1] = o - B + n
            hv["pf_ctx"][j] = a + n
            max_q = max(max_q, n)
            if c.samples:
                rows.append(o + n - 1)
                sampling.append(c)
            o += n
        n_rows = len(rows)
        hv["rows"][:n_rows] = rows
        k = 0
        for r, c in enumerate(sampling):
            s = c.seq
            # padded graph rows sit between decodes and prefill rows: map sampling index -> row slot
            slot_r = r if r < n_dec else (B - n_dec) + r
            hv["temps_f"][slot_r] = s.params.temperature
            hv["top_p_f"][slot_r] = s.params.top_p
            hv["seeds_q"][slot_r] = s.seed
            hv["steps"][slot_r] = s.num_generated + (1 if chained else 0)  # the in-flight step adds one token
            k += 1
        for i in range(n_dec, B):
            hv["temps_f"][i] = 0.0
            hv["top_p_f"][i] = 1.0

        # ---- host -> device (pinned, async on the compute stream) --------------------------------
        self._h2d("ids", T)
        self._h2d("pos", T)
        self._h2d("slots", T)
        lo, hi = self._meta_lo, self._meta_hi
        self.dbuf[lo:hi].copy_(self.hbuf[lo:hi], non_blocking=True)
        self.h2d_bytes += 4 * (hi - lo)
        if n_dec:
            self._h2d("dec_bt", n_dec * self.MB)
        if n_pf:
            self._h2d("pf_bt", n_pf * self.MB)
        if chained:
            o_ids, _ = self._off["ids"]
            self.dbuf[o_ids:o_ids + n_dec].copy_(self.out_ids[:n_dec], non_blocking=True)
        self._h2d_done.record()

        splits = self.decode_splits(B, max_len)
        whole_rows = self.tp > 1 and any(c.seq.params.top_p < 1.0 and c.seq.params.temperature > 0.0 for c in sampling)
        ev0 = None
        if self.step_events is not None:
            ev0 = torch.cuda.Event(enable_timing=True)
            ev0.record()
        if use_graph:
            key = (B, splits, whole_rows)
            g = self._graphs.get(key)
            if g is None:
                g = self._capture(B, splits, whole_rows)
            g.replay()
            self.launches_last_step = self._graph_launches[key]
        else:
            self.launches_last_step = self._forward(T, B, splits, n_pf, max_q, n_rows, whole_rows)
        self.kernel_launches += self.launches_last_step
        if ev0 is not None:
            ev1 = torch.cuda.Event(enable_timing=True)
            ev1.record()
            self.step_events.append((ev0, ev1, n_dec, T_real - n_dec, sum(c.start + 1 for c in decs)))

        ev = None
        buf = self._out_flip
        if n_rows:
            self._out_flip ^= 1
            self.out_host[buf][:n_rows].copy_(self.out_ids[:n_rows], non_blocking=True)
            self.d2h_bytes += 4 * n_rows
            ev = self._out_events[buf]
            ev.record()
        return (sampling, n_dec, B, buf, ev)

    def _capture(self, B: int, splits: int, whole_rows: bool = False) -> torch.cuda.CUDAGraph:
        """Capture one decode step for batch bucket B. Inputs already sit in dbuf; the warm-up run
        writes the same K/V slots and logits the replay will write again, so it is harmless."""
        torch.cuda.current_stream().synchronize()
        s = torch.cuda.Stream(device=self.dev)
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            self._forward(B, B, splits, 0, 0, B, whole_rows)  # warm-up: sets func attributes outside capture
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize(self.dev)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s, capture_error_mode="thread_local"):
            n = self._forward(B, B, splits, 0, 0, B, whole_rows)
        self._graphs[(B, splits, whole_rows)] = g
        self._graph_launches[(B, splits, whole_rows)] = n
        return g

    def close(self) -> None:
        """Drop captured graphs (they pin NCCL work under TP: destroying the process group while they
        are alive dead-locks at exit) and drain the stream."""
        self._graphs.clear()
        self._graph_launches.clear()
        torch.cuda.synchronize(self.dev)

    # ---- capacity facts for the server's gauges ---------------------------------------------------
    def kv_cache_info(self) -> Tuple[int, int]:
        return self.num_blocks, self.bs
