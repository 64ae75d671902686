// Paged-KV decode attention (one query token per sequence), split along the KV length.
//
// Arithmetic follows eager_attention_forward (transformers/models/llama/modeling_llama.py:199-222):
// scores = q.k * head_dim**-0.5, softmax in fp32, out = P.V, with repeat_kv (:187-196) realised by
// letting the G = Hq/Hkv query heads of one KV head share every K/V row that is read.
//
// Decode attention is HBM-bound integer-indexed byte streaming. grid = (Hkv, B, splits), ONE WARP per
// CTA: the G query heads of a KV head are rows of a 16-row mma.sync m16n8k16 tile (rows >= G are zero
// padding), so every K/V row is read once for all of them and a warp spends ~8 instructions per 512-byte
// row (a CUDA-core version spent ~90: the chip's whole issue rate at 6.5 TB/s; ncu showed 65 % issue
// slots at 7 % DRAM, profiles/r01b). K/V pages are staged with cp.async into a 3-deep ring of XOR-swizzled
// 16-token tiles (zero-filled past the split end) and read back with ldmatrix; the online softmax state
// lives in fp32 registers. Requires block_size % 16 == 0 (a tile never straddles a page) and G <= 8.
//
// Per-split partials (max, sum, out) go to caller scratch and attn_decode_reduce_kernel merges them.
//
// Measured and dropped (in-graph numbers in DESIGN.md section 4, profiles/r02e_*): a CUDA-core kernel (3.7 TB/s
// ceiling), a thread-block-cluster merge of the splits, RoPE + KV write in the qkv GEMM's epilogue, and a one-launch
// form of the whole decode attention path (RoPE on q in registers, the owner of the last tile writes the new K/V
// row, the split that arrives last merges): 14.0 us per layer against 10.9 us for rope_kv_write -> this kernel ->
// reduce under programmatic dependent launch, whose three short latency chains overlap each other's prefetch.
#include "common.cuh"

#include <cstdlib>
#include <mutex>

#include "../../include/atx_b200.h"

namespace atx {

// exp2(a - b) with the convention that an empty state (a == -inf) contributes nothing
__device__ __forceinline__ float safe_exp2_diff(float a, float b) {
    return a == -INFINITY ? 0.f : exp2f(a - b);
}

__device__ __forceinline__ void dec_cp_async16(void* smem_dst, const void* gsrc, bool valid) {
    const uint32_t sz = valid ? 16u : 0u;  // src-size 0 => zero fill: stale page rows (even NaN) never reach P.V
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(sz)
                 : "memory");
}
__device__ __forceinline__ void dec_ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
                 : "r"(addr));
}
__device__ __forceinline__ void dec_ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
                 : "r"(addr));
}
__device__ __forceinline__ void dec_mma(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                        uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, "
        "{%0, %1, %2, %3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

constexpr int kMmaTile = 16;    // tokens per tile
constexpr int kMmaStages = 3;

template <int D>
__global__ void __launch_bounds__(32)
attn_decode_mma_kernel(const __nv_bfloat16* __restrict__ q, const __nv_bfloat16* __restrict__ k_cache,
                       const __nv_bfloat16* __restrict__ v_cache, const int* __restrict__ block_table,
                       const int* __restrict__ seq_lens, __nv_bfloat16* __restrict__ out, float* __restrict__ ws_o,
                       float* __restrict__ ws_ml, int Hq, int Hkv, int block_size, int max_blocks, int num_splits,
                       float scale_log2) {
    constexpr int CH = D / 8;                     // 16-byte chunks per row
    constexpr int TILE_BYTES = kMmaTile * D * 2;  // one K (or V) tile
    __shared__ __align__(128) uint8_t smem[kMmaStages * 2 * TILE_BYTES];

    pdl_launch_dependents();
    const int g = blockIdx.x, b = blockIdx.y, sp = blockIdx.z;
    const int lane = threadIdx.x;
    const int gid = lane >> 2, tig = lane & 3;
    const int G = Hq / Hkv;
    const int L = seq_lens[b];
    int chunk = (L + num_splits - 1) / num_splits;
    chunk = (chunk + kMmaTile - 1) / kMmaTile * kMmaTile;
    const int begin = sp * chunk;
    const int end = min(L, begin + chunk);
    const int n_tiles = end > begin ? (end - begin + kMmaTile - 1) / kMmaTile : 0;
    const int* bt = block_table + static_cast<long long>(b) * max_blocks;

    auto issue_tile = [&](int t) {
        // a 16-token tile lies inside one page: one block-table lookup, rows are contiguous
        const int tok0 = begin + t * kMmaTile;
        const int page = __ldg(bt + tok0 / block_size);
        const long long base = ((static_cast<long long>(page) * Hkv + g) * block_size + tok0 % block_size) * D;
        uint8_t* sk = smem + (t % kMmaStages) * 2 * TILE_BYTES;
        uint8_t* sv = sk + TILE_BYTES;
#pragma unroll
        for (int i = lane; i < kMmaTile * CH; i += 32) {
            const int r = i / CH, c = i % CH;
            const uint32_t so = static_cast<uint32_t>(r * (D * 2) + ((c ^ (r & 7)) << 4));
            const bool valid = tok0 + r < end;
            dec_cp_async16(sk + so, k_cache + base + r * D + c * 8, valid);
            dec_cp_async16(sv + so, v_cache + base + r * D + c * 8, valid);
        }
    };
    // Sequence lengths, block tables and every K/V row except the one this step writes (always in the last
    // tile of the sequence) do not depend on the upstream kernel: those tiles are requested BEFORE
    // griddepcontrol.wait. Commit groups stay in tile order.
    const int fresh_tile = (end == L && n_tiles > 0) ? n_tiles - 1 : -1;
    int t_pre = 0;
    for (; t_pre < kMmaStages - 1; ++t_pre) {
        if (t_pre >= n_tiles || t_pre == fresh_tile) break;
        issue_tile(t_pre);
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    pdl_wait();

    for (int t = t_pre; t < kMmaStages - 1; ++t) {
        if (t < n_tiles) issue_tile(t);
        asm volatile("cp.async.commit_group;" ::: "memory");
    }

    // Q as the A operand: row gid = query head gid of this KV head (zero rows beyond G)
    uint32_t qa[D / 16][2];
    {
        const uint32_t* qrow =
            reinterpret_cast<const uint32_t*>(q + (static_cast<long long>(b) * Hq + g * G + min(gid, G - 1)) * D);
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
            qa[kk][0] = gid < G ? __ldg(qrow + kk * 8 + tig) : 0u;
            qa[kk][1] = gid < G ? __ldg(qrow + kk * 8 + 4 + tig) : 0u;
        }
    }
    float o_acc[D / 8][4];
#pragma unroll
    for (int i = 0; i < D / 8; ++i) o_acc[i][0] = o_acc[i][1] = o_acc[i][2] = o_acc[i][3] = 0.f;
    float m_run = -INFINITY, l_run = 0.f;

    for (int t = 0; t < n_tiles; ++t) {
        if (t + kMmaStages - 1 < n_tiles) issue_tile(t + kMmaStages - 1);
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group %0;" ::"n"(kMmaStages - 1) : "memory");
        __syncwarp();
        const uint8_t* sk = smem + (t % kMmaStages) * 2 * TILE_BYTES;
        const uint8_t* sv = sk + TILE_BYTES;

        float sc[2][4];
#pragma unroll
        for (int nb = 0; nb < 2; ++nb) sc[nb][0] = sc[nb][1] = sc[nb][2] = sc[nb][3] = 0.f;
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
            uint32_t b0, b1, b2, b3;
            const int r = (lane & 7) + (lane >> 4) * 8;
            const int c = kk * 2 + ((lane >> 3) & 1);
            dec_ldsm_x4(smem_u32(sk + r * (D * 2) + ((c ^ (r & 7)) << 4)), b0, b1, b2, b3);
            dec_mma(sc[0], qa[kk][0], 0u, qa[kk][1], 0u, b0, b1);
            dec_mma(sc[1], qa[kk][0], 0u, qa[kk][1], 0u, b2, b3);
        }
        // rows gid (valid heads); mask tokens past the end of this split
        const int tok0 = begin + t * kMmaTile;
        float mx = -INFINITY;
#pragma unroll
        for (int nb = 0; nb < 2; ++nb) {
            const int col = tok0 + nb * 8 + 2 * tig;
            if (col >= end) sc[nb][0] = -INFINITY;
            if (col + 1 >= end) sc[nb][1] = -INFINITY;
            mx = fmaxf(mx, fmaxf(sc[nb][0], sc[nb][1]));
        }
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
        const float mn = fmaxf(m_run, mx);  // finite: the first token of a tile is always < end
        const float mref = mn * scale_log2;
        const float corr = exp2f(m_run * scale_log2 - mref);
        m_run = mn;
        const float p00 = exp2f(sc[0][0] * scale_log2 - mref), p01 = exp2f(sc[0][1] * scale_log2 - mref);
        const float p10 = exp2f(sc[1][0] * scale_log2 - mref), p11 = exp2f(sc[1][1] * scale_log2 - mref);
        l_run = l_run * corr + (p00 + p01) + (p10 + p11);
        const uint32_t pa0 = pack_bf16(p00, p01), pa2 = pack_bf16(p10, p11);
#pragma unroll
        for (int db = 0; db < D / 8; ++db) {
            o_acc[db][0] *= corr;
            o_acc[db][1] *= corr;
        }
#pragma unroll
        for (int db2 = 0; db2 < D / 16; ++db2) {
            uint32_t b0, b1, b2, b3;
            const int r = (lane & 7) + ((lane >> 3) & 1) * 8;
            const int c = db2 * 2 + (lane >> 4);
            dec_ldsm_x4_t(smem_u32(sv + r * (D * 2) + ((c ^ (r & 7)) << 4)), b0, b1, b2, b3);
            dec_mma(o_acc[2 * db2], pa0, 0u, pa2, 0u, b0, b1);
            dec_mma(o_acc[2 * db2 + 1], pa0, 0u, pa2, 0u, b2, b3);
        }
        __syncwarp();  // all lanes done reading this stage before it is refilled
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");

    l_run += __shfl_xor_sync(0xffffffffu, l_run, 1);
    l_run += __shfl_xor_sync(0xffffffffu, l_run, 2);
    if (gid < G) {
        const long long h = static_cast<long long>(b) * Hq + g * G + gid;
        if (num_splits == 1) {
            const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
#pragma unroll
            for (int db = 0; db < D / 8; ++db)
                *reinterpret_cast<uint32_t*>(out + h * D + db * 8 + 2 * tig) =
                    pack_bf16(o_acc[db][0] * inv, o_acc[db][1] * inv);
        } else {
            float* po = ws_o + (h * num_splits + sp) * D;
#pragma unroll
            for (int db = 0; db < D / 8; ++db)
                *reinterpret_cast<float2*>(po + db * 8 + 2 * tig) = make_float2(o_acc[db][0], o_acc[db][1]);
            if (tig == 0) {
                ws_ml[(h * num_splits + sp) * 2] = m_run == -INFINITY ? -INFINITY : m_run * scale_log2;
                ws_ml[(h * num_splits + sp) * 2 + 1] = l_run;
            }
        }
    }
}

// out[h, :] = sum_s o_s * 2^(m_s - M) / sum_s l_s * 2^(m_s - M)
__global__ void attn_decode_reduce_kernel(const float* __restrict__ ws_o, const float* __restrict__ ws_ml,
                                          __nv_bfloat16* __restrict__ out, int D, int num_splits) {
    pdl_launch_dependents();
    pdl_wait();
    const long long h = blockIdx.x;
    const float* ml = ws_ml + h * num_splits * 2;
    float M = -INFINITY;
    for (int s = 0; s < num_splits; ++s) M = fmaxf(M, ml[2 * s]);
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
        float o = 0.f, ls = 0.f;
        for (int s = 0; s < num_splits; ++s) {
            const float f = safe_exp2_diff(ml[2 * s], M);
            o += ws_o[(h * num_splits + s) * D + d] * f;
            ls += ml[2 * s + 1] * f;
        }
        out[h * D + d] = __float2bfloat16_rn(ls > 0.f ? o / ls : 0.f);
    }
}

template <int D>
static int launch_decode(const void* q, const void* kc, const void* vc, const int* bt, const int* sl, void* out,
                         float* ws_o, float* ws_ml, int B, int Hq, int Hkv, int block_size, int max_blocks,
                         int num_splits, float scale, int flags, cudaStream_t stream) {
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = (flags & ATX_FLAG_PDL) ? 1 : 0;
    cfg.gridDim = dim3(static_cast<unsigned>(Hkv), static_cast<unsigned>(B), static_cast<unsigned>(num_splits));
    cfg.blockDim = dim3(32);
    cfg.stream = stream;
    ATX_CUDA(cudaLaunchKernelEx(&cfg, attn_decode_mma_kernel<D>, reinterpret_cast<const __nv_bfloat16*>(q),
                                reinterpret_cast<const __nv_bfloat16*>(kc), reinterpret_cast<const __nv_bfloat16*>(vc),
                                bt, sl, reinterpret_cast<__nv_bfloat16*>(out), ws_o, ws_ml, Hq, Hkv, block_size,
                                max_blocks, num_splits, scale * 1.4426950408889634f));
    if (num_splits > 1) {
        cfg.gridDim = dim3(static_cast<unsigned>(B * Hq));
        cfg.blockDim = dim3(static_cast<unsigned>(D));
        ATX_CUDA(cudaLaunchKernelEx(&cfg, attn_decode_reduce_kernel, static_cast<const float*>(ws_o),
                                    static_cast<const float*>(ws_ml), reinterpret_cast<__nv_bfloat16*>(out), D,
                                    num_splits));
    }
    return ATX_OK;
}

static int check_decode_shape(const char* who, int B, int Hq, int Hkv, int D, int block_size, int max_blocks,
                              int num_splits) {
    ATX_REQUIRE(B > 0 && Hq > 0 && Hkv > 0 && Hq % Hkv == 0 && block_size > 0 && max_blocks > 0 && num_splits > 0,
                "%s: bad shape B=%d Hq=%d Hkv=%d block_size=%d max_blocks=%d splits=%d", who, B, Hq, Hkv, block_size,
                max_blocks, num_splits);
    ATX_REQUIRE(num_splits <= 65535 && B <= 65535, "%s: B and num_splits must be <= 65535", who);
    if ((D != 64 && D != 128) || Hq / Hkv > 8 || block_size % kMmaTile != 0) {
        set_error("%s: unsupported head_dim=%d / group=%d / block_size=%d (head_dim 64|128, at most 8 query heads per "
                  "KV head, pages of a multiple of 16 tokens)", who, D, Hq / Hkv, block_size);
        return ATX_ERR_UNSUPPORTED;
    }
    return ATX_OK;
}

}  // namespace atx

using namespace atx;

extern "C" {

long long atx_paged_attn_decode_workspace_bytes(int B, int Hq, int D, int num_splits) {
    if (num_splits <= 1) return 0;
    return static_cast<long long>(B) * Hq * num_splits * (D + 2) * 4;
}

int atx_paged_attn_decode(const void* q, const void* kc, const void* vc, const int* bt, const int* sl,
                          void* out, void* workspace, long long workspace_bytes, int B, int Hq, int Hkv, int D,
                          int block_size, int max_blocks, int num_splits, float scale, int flags,
                          atx_stream_t stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    ATX_REQUIRE(q && kc && vc && bt && sl && out, "atx_paged_attn_decode: null pointer");
    int rc = check_decode_shape("atx_paged_attn_decode", B, Hq, Hkv, D, block_size, max_blocks, num_splits);
    if (rc != ATX_OK) return rc;
    float* ws_o = nullptr;
    float* ws_ml = nullptr;
    if (num_splits > 1) {
        const long long need = atx_paged_attn_decode_workspace_bytes(B, Hq, D, num_splits);
        ATX_REQUIRE(workspace && workspace_bytes >= need,
                    "atx_paged_attn_decode: workspace of %lld bytes needed, %lld given", need, workspace_bytes);
        ws_o = static_cast<float*>(workspace);
        ws_ml = ws_o + static_cast<long long>(B) * Hq * num_splits * D;
    }
    if (D == 128)
        return launch_decode<128>(q, kc, vc, bt, sl, out, ws_o, ws_ml, B, Hq, Hkv, block_size, max_blocks,
                                         num_splits, scale, flags, stream);
    return launch_decode<64>(q, kc, vc, bt, sl, out, ws_o, ws_ml, B, Hq, Hkv, block_size, max_blocks,
                                    num_splits, scale, flags, stream);
}

// ---- decode attention without any allocation -----------------------------------------------------
// Some hosts of this library can hand us pointers but cannot obtain device memory themselves: a
// server that has already given every free byte of HBM to its KV pool, a process running under a
// fixed memory reservation where a late cudaMalloc fails, a request path that must not take
// cudaMalloc's device-wide synchronisation, or a foreign-language binding (cgo / JNI / N-API) that
// has no CUDA runtime of its own to allocate the split-KV scratch with. For them the partial
// results of the KV splits live in the array below, which is part of the library image: every GPU
// that loads libatx_b200.so holds these 8 MiB until the process exits.
constexpr long long kStaticScratchBytes = 8ll << 20;
static __device__ __align__(256) float g_decode_scratch[kStaticScratchBytes / 4];

// `g_decode_scratch` in host code is only a handle for the runtime, not an address a kernel may
// dereference, so the device address is looked up once per device (with that device current).
static float* static_scratch_for_current_device(int* rc) {
    static std::mutex mu;
    static float* ptr[64] = {nullptr};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) {
        *rc = cuda_fail(e, "cudaGetDevice");
        return nullptr;
    }
    if (dev < 0 || dev >= 64) {
        set_error("atx_paged_attn_decode_static: device ordinal %d is outside the supported range [0, 64)", dev);
        *rc = ATX_ERR_LIMIT;
        return nullptr;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (ptr[dev] == nullptr) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_decode_scratch);
        if (e != cudaSuccess) {
            *rc = cuda_fail(e, "cudaGetSymbolAddress(g_decode_scratch)");
            return nullptr;
        }
        ptr[dev] = static_cast<float*>(p);
    }
    *rc = ATX_OK;
    return ptr[dev];
}

long long atx_paged_attn_decode_static_capacity(void) { return kStaticScratchBytes; }

int atx_paged_attn_decode_static(const void* q, const void* kc, const void* vc, const int* bt, const int* sl,
                                 void* out, int B, int Hq, int Hkv, int D, int block_size, int max_blocks,
                                 int num_splits, float scale, int flags, atx_stream_t stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    ATX_REQUIRE(q && kc && vc && bt && sl && out, "atx_paged_attn_decode_static: null pointer");
    int shape_rc = check_decode_shape("atx_paged_attn_decode_static", B, Hq, Hkv, D, block_size, max_blocks, num_splits);
    if (shape_rc != ATX_OK) return shape_rc;
    const long long need = atx_paged_attn_decode_workspace_bytes(B, Hq, D, num_splits);
    if (need > kStaticScratchBytes) {
        set_error("atx_paged_attn_decode_static: B=%d Hq=%d D=%d num_splits=%d needs %lld bytes of split scratch, the "
                  "built-in array holds %lld; use fewer splits / a smaller batch, or atx_paged_attn_decode with "
                  "caller-provided scratch",
                  B, Hq, D, num_splits, need, kStaticScratchBytes);
        return ATX_ERR_LIMIT;
    }
    float* ws_o = nullptr;
    float* ws_ml = nullptr;
    if (num_splits > 1) {
        int rc = ATX_OK;
        float* base = static_scratch_for_current_device(&rc);
        if (base == nullptr) return rc;
        // Nothing in the scratch has to start at zero: every (head, split) slot that the merge
        // kernel reads is rewritten by this call's own attention kernel first.
        ws_o = base;
        ws_ml = ws_o + static_cast<long long>(B) * Hq * num_splits * D;
    }
    if (D == 128)
        return launch_decode<128>(q, kc, vc, bt, sl, out, ws_o, ws_ml, B, Hq, Hkv, block_size, max_blocks,
                                         num_splits, scale, flags, stream);
    return launch_decode<64>(q, kc, vc, bt, sl, out, ws_o, ws_ml, B, Hq, Hkv, block_size, max_blocks,
                                    num_splits, scale, flags, stream);
}

}  // extern "C"
