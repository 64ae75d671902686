"""ctypes binding of libatx_b200.so (include/atx_b200.h). No torch types cross the ABI: tensors are
passed as raw device pointers + sizes; torch only owns the memory and the stream.

There is NO fallback: if the library is missing or a call fails, this raises.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path
from typing import Optional

import torch

_PKG_ROOT = Path(__file__).resolve().parent.parent
_LIB_PATH = _PKG_ROOT / "lib" / "libatx_b200.so"

GEMM_STORE_F32 = 0
GEMM_ADD_F32 = 1
GEMM_STORE_BF16 = 2
GEMM_SWIGLU_BF16 = 3
FLAG_PDL = 1
FLAG_ORDERED = 2   # atx_gemm_run ADD_F32, T <= 64: fixed-order split-K (bit-reproducible)
FLAG_ADD_BF16 = 4  # atx_rmsnorm: the `add` buffer is bf16

EXPORTS = [
    "atx_last_error", "atx_version", "atx_device_info",
    "atx_gemm_plan_create", "atx_gemm_plan_destroy", "atx_gemm_run", "atx_gemm_auto_splits",
    "atx_embed_gather", "atx_rmsnorm", "atx_rope_kv_write", "atx_silu_mul",
    "atx_reduce_scatter_rmsnorm_all_gather", "atx_allreduce_add_rmsnorm_ll",
    "atx_paged_attn_decode_workspace_bytes", "atx_paged_attn_decode", "atx_paged_attn_prefill",
    "atx_paged_attn_decode_static", "atx_paged_attn_decode_static_capacity",
    "atx_sample", "atx_sample_shard", "atx_sample_merge",
]


class AtxError(RuntimeError):
    pass


_lib: Optional[C.CDLL] = None


def lib_path() -> Path:
    return Path(os.environ.get("ATX_B200_LIB", str(_LIB_PATH)))


def load() -> C.CDLL:
    """Load the shared library (building it first if nvcc and the sources are here)."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not path.exists():
        import importlib.util

        spec = importlib.util.spec_from_file_location("atx_build", _PKG_ROOT / "build.py")
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)  # type: ignore[union-attr]
        mod.build_lib()
    if not path.exists():
        raise AtxError(f"libatx_b200.so not found at {path}; run `python agentic-traffic-testing_b200/build.py`")
    lib = C.CDLL(str(path))
    vp, i32, i64, f32 = C.c_void_p, C.c_int, C.c_longlong, C.c_float
    lib.atx_last_error.restype = C.c_char_p
    lib.atx_last_error.argtypes = []
    lib.atx_version.restype = i32
    lib.atx_device_info.argtypes = [C.POINTER(i32)] * 3
    lib.atx_gemm_plan_create.argtypes = [C.POINTER(vp), vp, i64, i64, vp, i64, i64]
    lib.atx_gemm_plan_destroy.argtypes = [vp]
    lib.atx_gemm_plan_destroy.restype = None
    lib.atx_gemm_run.argtypes = [vp, vp, i64, i32, i32, i32, i32, vp]
    lib.atx_gemm_auto_splits.argtypes = [i64, i64, i32]
    lib.atx_embed_gather.argtypes = [vp, vp, vp, i32, i32, i32, vp]
    lib.atx_rmsnorm.argtypes = [vp, vp, vp, vp, vp, i32, i32, f32, vp, i64, i32, vp]
    lib.atx_reduce_scatter_rmsnorm_all_gather.argtypes = [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), vp, i32, vp,
                                                          i32, i32, vp, vp, i32, i32, f32, vp, i64, i32, vp]
    lib.atx_allreduce_add_rmsnorm_ll.argtypes = [vp, vp, C.POINTER(vp), C.POINTER(vp), vp, i32, vp, i64, i32, i32, i32,
                                                 vp, vp, i32, i32, f32, vp, i64, i32, vp]
    lib.atx_rope_kv_write.argtypes = [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp]
    lib.atx_silu_mul.argtypes = [vp, vp, i32, i32, i32, vp]
    lib.atx_paged_attn_decode_workspace_bytes.argtypes = [i32, i32, i32, i32]
    lib.atx_paged_attn_decode_workspace_bytes.restype = i64
    lib.atx_paged_attn_decode.argtypes = [vp, vp, vp, vp, vp, vp, vp, i64, i32, i32, i32, i32, i32, i32, i32, f32,
                                          i32, vp]
    lib.atx_paged_attn_decode_static.argtypes = [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, f32, i32, vp]
    lib.atx_paged_attn_decode_static_capacity.argtypes = []
    lib.atx_paged_attn_decode_static_capacity.restype = i64
    lib.atx_paged_attn_prefill.argtypes = [vp, vp, vp, vp, vp, vp, vp, i32, i32, i64, i64, i32, i32, i32, i32, i32, f32,
                                           i32, vp]
    lib.atx_sample.argtypes = [vp, i64, vp, vp, vp, vp, vp, i32, i32, i32, vp]
    lib.atx_sample_shard.argtypes = [vp, i64, i32, vp, vp, vp, vp, i32, i32, i32, vp]
    lib.atx_sample_merge.argtypes = [vp, i32, i32, vp, i32, vp]
    for name in EXPORTS:  # fail at load time, not at first use, if the build is stale
        getattr(lib, name)
    _lib = lib
    return lib


def _check(rc: int) -> None:
    if rc != 0:
        raise AtxError(f"atx error {rc}: {load().atx_last_error().decode()}")


def _p(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _dev(*ts: Optional[torch.Tensor]) -> None:
    for t in ts:
        if t is not None and (not t.is_cuda or not t.is_contiguous()):
            raise AtxError("atx kernels need contiguous CUDA tensors")


class GemmPlan:
    """out[t, n] (+)= X[t, :] . W[n, :]  — holds the TMA descriptors for one (W, X buffer) pair."""

    def __init__(self, w: torch.Tensor, x_buf: torch.Tensor):
        _dev(w, x_buf)
        assert w.dtype == torch.bfloat16 and x_buf.dtype == torch.bfloat16 and w.dim() == 2 and x_buf.dim() == 2
        assert w.shape[1] == x_buf.shape[1]
        self.w, self.x = w, x_buf  # keep alive
        self.N, self.K = int(w.shape[0]), int(w.shape[1])
        self.t_cap = int(x_buf.shape[0])
        h = C.c_void_p()
        _check(load().atx_gemm_plan_create(C.byref(h), w.data_ptr(), self.N, self.K, x_buf.data_ptr(), self.t_cap,
                                           x_buf.stride(0)))
        self._h = h

    def run(self, out: torch.Tensor, T: int, mode: int, splits: int = 0, flags: int = 0) -> None:
        _dev(out)
        want = torch.bfloat16 if mode in (GEMM_STORE_BF16, GEMM_SWIGLU_BF16) else torch.float32
        cols = self.N // 2 if mode == GEMM_SWIGLU_BF16 else self.N
        assert out.dtype == want and out.dim() == 2 and out.shape[1] >= cols and out.shape[0] >= T
        _check(load().atx_gemm_run(self._h, out.data_ptr(), out.stride(0), T, mode, splits, flags, _stream()))

    def __del__(self):
        try:
            if getattr(self, "_h", None) and _lib is not None:
                _lib.atx_gemm_plan_destroy(self._h)
        except Exception:
            pass


def gemm_auto_splits(N: int, K: int, T: int) -> int:
    return int(load().atx_gemm_auto_splits(N, K, T))


def embed_gather(ids: torch.Tensor, table: torch.Tensor, out: torch.Tensor, T: int) -> None:
    _dev(ids, table, out)
    assert ids.dtype == torch.int32 and table.dtype == torch.bfloat16 and out.dtype == torch.float32
    _check(load().atx_embed_gather(ids.data_ptr(), table.data_ptr(), out.data_ptr(), T, table.shape[1],
                                   table.shape[0], _stream()))


def rmsnorm(x: torch.Tensor, w: torch.Tensor, out: torch.Tensor, T: int, eps: float,
            rows: Optional[torch.Tensor] = None, zero: Optional[torch.Tensor] = None, zero_bytes: int = 0,
            flags: int = 0, add: Optional[torch.Tensor] = None) -> None:
    _dev(x, w, out, rows, zero, add)
    assert add is None or (add.shape[-1] == x.shape[-1] and
                           add.dtype == (torch.bfloat16 if flags & FLAG_ADD_BF16 else torch.float32))
    assert x.dtype == torch.float32 and w.dtype == torch.bfloat16 and out.dtype == torch.bfloat16
    if zero is not None and zero_bytes == 0:
        zero_bytes = zero.numel() * zero.element_size()
    _check(load().atx_rmsnorm(x.data_ptr(), _p(add), _p(rows), w.data_ptr(), out.data_ptr(), T, x.shape[-1], eps, _p(zero),
                              zero_bytes, flags, _stream()))


def rope_kv_write(qkv: torch.Tensor, positions: torch.Tensor, slots: torch.Tensor, cos_sin: torch.Tensor,
                  q_out: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, T: int, Hq: int, Hkv: int,
                  D: int, block_size: int, flags: int = 0) -> None:
    _dev(qkv, positions, slots, cos_sin, q_out, k_cache, v_cache)
    assert qkv.dtype == torch.float32 and positions.dtype == torch.int32 and slots.dtype == torch.int32
    assert cos_sin.dtype == torch.float32 and q_out.dtype == torch.bfloat16 and k_cache.dtype == torch.bfloat16
    _check(load().atx_rope_kv_write(qkv.data_ptr(), positions.data_ptr(), slots.data_ptr(), cos_sin.data_ptr(),
                                    q_out.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(), T, Hq, Hkv, D,
                                    block_size, flags, _stream()))


def silu_mul(gu: torch.Tensor, act: torch.Tensor, T: int, I: int, flags: int = 0) -> None:
    _dev(gu, act)
    assert gu.dtype == torch.float32 and act.dtype == torch.bfloat16
    _check(load().atx_silu_mul(gu.data_ptr(), act.data_ptr(), T, I, flags, _stream()))


def decode_workspace_bytes(B: int, Hq: int, D: int, splits: int) -> int:
    return int(load().atx_paged_attn_decode_workspace_bytes(B, Hq, D, splits))


def paged_attn_decode(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, block_table: torch.Tensor,
                      seq_lens: torch.Tensor, out: torch.Tensor, workspace: Optional[torch.Tensor], B: int, Hq: int,
                      Hkv: int, D: int, block_size: int, num_splits: int, scale: float, flags: int = 0) -> None:
    _dev(q, k_cache, v_cache, block_table, seq_lens, out, workspace)
    assert q.dtype == torch.bfloat16 and out.dtype == torch.bfloat16 and block_table.dtype == torch.int32
    assert seq_lens.dtype == torch.int32
    wsb = 0 if workspace is None else workspace.numel() * workspace.element_size()
    _check(load().atx_paged_attn_decode(q.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(), block_table.data_ptr(),
                                        seq_lens.data_ptr(), out.data_ptr(), _p(workspace), wsb, B, Hq, Hkv, D,
                                        block_size, block_table.shape[1], num_splits, scale, flags, _stream()))


def paged_attn_decode_static(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, block_table: torch.Tensor,
                             seq_lens: torch.Tensor, out: torch.Tensor, B: int, Hq: int, Hkv: int, D: int,
                             block_size: int, num_splits: int, scale: float, flags: int = 0) -> None:
    """Decode attention whose split scratch is the library's built-in 8 MiB array: for callers that
    cannot allocate device memory (see include/atx_b200.h). Calls on one device must not overlap."""
    _dev(q, k_cache, v_cache, block_table, seq_lens, out)
    assert q.dtype == torch.bfloat16 and out.dtype == torch.bfloat16 and block_table.dtype == torch.int32
    assert seq_lens.dtype == torch.int32
    _check(load().atx_paged_attn_decode_static(q.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(),
                                               block_table.data_ptr(), seq_lens.data_ptr(), out.data_ptr(), B, Hq,
                                               Hkv, D, block_size, block_table.shape[1], num_splits, scale, flags,
                                               _stream()))


def decode_static_capacity() -> int:
    return int(load().atx_paged_attn_decode_static_capacity())


def paged_attn_prefill(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, block_table: torch.Tensor,
                       cu_q: torch.Tensor, ctx_lens: torch.Tensor, out: torch.Tensor, S: int, max_q_len: int, Hq: int,
                       Hkv: int, D: int, block_size: int, scale: float, flags: int = 0) -> None:
    _dev(q, k_cache, v_cache, block_table, cu_q, ctx_lens, out)
    assert q.dtype == torch.bfloat16 and out.dtype == torch.bfloat16 and block_table.dtype == torch.int32
    assert cu_q.dtype == torch.int32 and ctx_lens.dtype == torch.int32
    _check(load().atx_paged_attn_prefill(q.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(), block_table.data_ptr(),
                                         cu_q.data_ptr(), ctx_lens.data_ptr(), out.data_ptr(), S, max_q_len,
                                         q.shape[0], k_cache.shape[0], Hq, Hkv, D, block_size, block_table.shape[1],
                                         scale, flags, _stream()))


def sample(logits: torch.Tensor, out_ids: torch.Tensor, B: int, V: int, temperature: Optional[torch.Tensor] = None,
           top_p: Optional[torch.Tensor] = None, seeds: Optional[torch.Tensor] = None,
           steps: Optional[torch.Tensor] = None, flags: int = 0) -> None:
    _dev(logits, out_ids, temperature, top_p, seeds, steps)
    assert logits.dtype == torch.float32 and out_ids.dtype == torch.int32
    assert temperature is None or temperature.dtype == torch.float32
    assert top_p is None or top_p.dtype == torch.float32
    assert seeds is None or seeds.dtype == torch.int64
    assert steps is None or steps.dtype == torch.int32
    _check(load().atx_sample(logits.data_ptr(), logits.stride(0), _p(temperature), _p(top_p), _p(seeds), _p(steps),
                             out_ids.data_ptr(), B, V, flags, _stream()))


def sample_shard(logits: torch.Tensor, idx_offset: int, pairs: torch.Tensor, B: int, V_local: int,
                 temperature: Optional[torch.Tensor] = None, seeds: Optional[torch.Tensor] = None,
                 steps: Optional[torch.Tensor] = None, flags: int = 0) -> None:
    _dev(logits, pairs, temperature, seeds, steps)
    assert logits.dtype == torch.float32 and pairs.dtype == torch.float32 and pairs.numel() >= 2 * B
    _check(load().atx_sample_shard(logits.data_ptr(), logits.stride(0), idx_offset, _p(temperature), _p(seeds), _p(steps),
                                   pairs.data_ptr(), B, V_local, flags, _stream()))


def sample_merge(pairs: torch.Tensor, world: int, B: int, out_ids: torch.Tensor, flags: int = 0) -> None:
    _dev(pairs, out_ids)
    assert pairs.dtype == torch.float32 and pairs.numel() >= 2 * world * B and out_ids.dtype == torch.int32
    _check(load().atx_sample_merge(pairs.data_ptr(), world, B, out_ids.data_ptr(), flags, _stream()))
