// Causal attention for prefill / chunked prefill over the paged KV cache.
//
// Arithmetic follows eager_attention_forward (transformers/models/llama/modeling_llama.py:199-222)
// with the additive causal mask, GQA by head sharing (repeat_kv :187-196) and fp32 softmax (:217).
// The chunk's own K/V are already in the pages (atx_rope_kv_write ran first), so every key is read
// through the block table: prefix-cache hits and chunked prefill need no special case.
//
// Tiling: one CTA = 64 query rows of one head (4 warps x 16 rows); KV is walked in 64-token tiles
// staged in XOR-swizzled shared memory with cp.async (double buffered, zero-filled past the
// context so that stale page contents can never reach the accumulators); QK^T and PV run on
// mma.sync m16n8k16 bf16 with fp32 accumulation and a flash-style online softmax in registers.
// TODO(round 2): move QK^T / PV to tcgen05 with S/P in TMEM for the 32k-token prefill config.
#include "common.cuh"

#include <cstdlib>

#include "../../include/atx_b200.h"

namespace atx {

constexpr int kPfBM = 64;
constexpr int kPfBN = 64;
constexpr int kPfThreads = 128;

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc, bool valid) {
    const uint32_t sz = valid ? 16u : 0u;  // src-size 0 => 16 bytes of zeros
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(sz)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
                 : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
                 : "r"(addr));
}
__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                               uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, "
        "{%0, %1, %2, %3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// Tile of `rows` x D bf16, rows of D*2 bytes, 16-byte chunks XOR-swizzled with (row & 7).
template <int D>
__device__ __forceinline__ uint32_t swz(int row, int chunk) {
    return static_cast<uint32_t>(row * (D * 2) + ((chunk ^ (row & 7)) << 4));
}

template <int D>
__global__ void __launch_bounds__(kPfThreads)
attn_prefill_kernel(const __nv_bfloat16* __restrict__ q, const __nv_bfloat16* __restrict__ k_cache,
                    const __nv_bfloat16* __restrict__ v_cache, const int* __restrict__ block_table,
                    const int* __restrict__ cu_q, const int* __restrict__ ctx_lens, __nv_bfloat16* __restrict__ out,
                    int Hq, int Hkv, int block_size, int max_blocks, float scale_log2) {
    constexpr int CH = D / 8;            // 16-byte chunks per row
    constexpr int TILE = kPfBN * D * 2;  // bytes of one K or V tile
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* sQ = smem;                  // 64 x D
    uint8_t* sK = smem + kPfBM * D * 2;  // 2 stages
    uint8_t* sV = sK + 2 * TILE;         // 2 stages

    pdl_launch_dependents();
    pdl_wait();
    const int s = blockIdx.z, h = blockIdx.y;
    const int q_begin = cu_q[s];
    const int n_q = cu_q[s + 1] - q_begin;
    const int m0 = blockIdx.x * kPfBM;
    if (m0 >= n_q) return;
    const int ctx = ctx_lens[s];
    const int off = ctx - n_q;  // query j sees keys [0, off + j]
    const int g = h / (Hq / Hkv);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int gid = lane >> 2, tig = lane & 3;
    const int* bt = block_table + static_cast<long long>(s) * max_blocks;

    const int kv_limit = min(ctx, off + m0 + kPfBM);  // keys any row of this tile may see
    const int n_tiles = (kv_limit + kPfBN - 1) / kPfBN;

    // ---- stage Q (once) and the first K/V tile
    for (int i = threadIdx.x; i < kPfBM * CH; i += kPfThreads) {
        const int r = i / CH, c = i % CH;
        const bool ok = m0 + r < n_q;
        const __nv_bfloat16* src = q + (static_cast<long long>(q_begin + (ok ? m0 + r : 0)) * Hq + h) * D + c * 8;
        cp_async16(sQ + swz<D>(r, c), src, ok);
    }
    auto load_kv = [&](int tile, int stage) {
        const int kv0 = tile * kPfBN;
        for (int i = threadIdx.x; i < kPfBN * CH; i += kPfThreads) {
            const int r = i / CH, c = i % CH;
            const int tok = kv0 + r;
            const bool ok = tok < ctx;
            const int tc = ok ? tok : 0;
            const int page = __ldg(bt + tc / block_size);
            const long long o =
                ((static_cast<long long>(page) * Hkv + g) * block_size + tc % block_size) * D + c * 8;
            cp_async16(sK + stage * TILE + swz<D>(r, c), k_cache + o, ok);
            cp_async16(sV + stage * TILE + swz<D>(r, c), v_cache + o, ok);
        }
    };
    load_kv(0, 0);
    cp_async_commit();

    float o_acc[D / 8][4];
#pragma unroll
    for (int i = 0; i < D / 8; ++i) o_acc[i][0] = o_acc[i][1] = o_acc[i][2] = o_acc[i][3] = 0.f;
    float m_run[2] = {-INFINITY, -INFINITY};  // rows gid and gid + 8
    float l_run[2] = {0.f, 0.f};
    uint32_t qa[D / 16][4];

    const int row_lo = m0 + warp * 16 + gid;  // query index within the sequence chunk
    const int row_hi = row_lo + 8;

    for (int tile = 0; tile < n_tiles; ++tile) {
        const int stage = tile & 1;
        if (tile + 1 < n_tiles) load_kv(tile + 1, stage ^ 1);
        cp_async_commit();
        cp_async_wait<1>();
        __syncthreads();

        if (tile == 0) {
            // Q fragments stay in registers for the whole KV walk
#pragma unroll
            for (int kk = 0; kk < D / 16; ++kk) {
                const int r = warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
                const int c = kk * 2 + (lane >> 4);
                ldsm_x4(smem_u32(sQ + swz<D>(r, c)), qa[kk][0], qa[kk][1], qa[kk][2], qa[kk][3]);
            }
        }

        // ---- S = Q K^T  (16 x 64 per warp)
        float sc[kPfBN / 8][4];
#pragma unroll
        for (int nb = 0; nb < kPfBN / 8; ++nb) sc[nb][0] = sc[nb][1] = sc[nb][2] = sc[nb][3] = 0.f;
        const uint8_t* kt = sK + stage * TILE;
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
#pragma unroll
            for (int nb2 = 0; nb2 < kPfBN / 16; ++nb2) {
                uint32_t b0, b1, b2, b3;
                const int r = nb2 * 16 + (lane & 7) + (lane >> 4) * 8;
                const int c = kk * 2 + ((lane >> 3) & 1);
                ldsm_x4(smem_u32(kt + swz<D>(r, c)), b0, b1, b2, b3);
                mma_bf16_16816(sc[2 * nb2], qa[kk][0], qa[kk][1], qa[kk][2], qa[kk][3], b0, b1);
                mma_bf16_16816(sc[2 * nb2 + 1], qa[kk][0], qa[kk][1], qa[kk][2], qa[kk][3], b2, b3);
            }
        }

        // ---- causal mask + online softmax (scores scaled by scale * log2 e inside exp2)
        const int kv0 = tile * kPfBN;
        const bool need_mask = kv0 + kPfBN - 1 > off + m0 + warp * 16;  // warp-uniform
        float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
        for (int nb = 0; nb < kPfBN / 8; ++nb) {
            if (need_mask) {
                const int col = kv0 + nb * 8 + 2 * tig;
                if (col > off + row_lo) sc[nb][0] = -INFINITY;
                if (col + 1 > off + row_lo) sc[nb][1] = -INFINITY;
                if (col > off + row_hi) sc[nb][2] = -INFINITY;
                if (col + 1 > off + row_hi) sc[nb][3] = -INFINITY;
            }
            mx[0] = fmaxf(mx[0], fmaxf(sc[nb][0], sc[nb][1]));
            mx[1] = fmaxf(mx[1], fmaxf(sc[nb][2], sc[nb][3]));
        }
        float corr[2], mref[2];
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            mx[i] = fmaxf(mx[i], __shfl_xor_sync(0xffffffffu, mx[i], 1));
            mx[i] = fmaxf(mx[i], __shfl_xor_sync(0xffffffffu, mx[i], 2));
            const float mn = fmaxf(m_run[i], mx[i]);
            mref[i] = mn == -INFINITY ? 0.f : mn * scale_log2;
            corr[i] = exp2f(m_run[i] * scale_log2 - mref[i]);  // first tile: exp2(-inf) = 0
            m_run[i] = mn;
        }
        float ls[2] = {0.f, 0.f};
        uint32_t pa[kPfBN / 16][4];
#pragma unroll
        for (int nb = 0; nb < kPfBN / 8; ++nb) {
            const float p0 = exp2f(sc[nb][0] * scale_log2 - mref[0]);
            const float p1 = exp2f(sc[nb][1] * scale_log2 - mref[0]);
            const float p2 = exp2f(sc[nb][2] * scale_log2 - mref[1]);
            const float p3 = exp2f(sc[nb][3] * scale_log2 - mref[1]);
            ls[0] += p0 + p1;
            ls[1] += p2 + p3;
            pa[nb >> 1][(nb & 1) * 2] = pack_bf16(p0, p1);
            pa[nb >> 1][(nb & 1) * 2 + 1] = pack_bf16(p2, p3);
        }
#pragma unroll
        for (int i = 0; i < 2; ++i) l_run[i] = l_run[i] * corr[i] + ls[i];
#pragma unroll
        for (int db = 0; db < D / 8; ++db) {
            o_acc[db][0] *= corr[0];
            o_acc[db][1] *= corr[0];
            o_acc[db][2] *= corr[1];
            o_acc[db][3] *= corr[1];
        }

        // ---- O += P V
        const uint8_t* vt = sV + stage * TILE;
#pragma unroll
        for (int kk = 0; kk < kPfBN / 16; ++kk) {
#pragma unroll
            for (int db2 = 0; db2 < D / 16; ++db2) {
                uint32_t b0, b1, b2, b3;
                const int r = kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
                const int c = db2 * 2 + (lane >> 4);
                ldsm_x4_t(smem_u32(vt + swz<D>(r, c)), b0, b1, b2, b3);
                mma_bf16_16816(o_acc[2 * db2], pa[kk][0], pa[kk][1], pa[kk][2], pa[kk][3], b0, b1);
                mma_bf16_16816(o_acc[2 * db2 + 1], pa[kk][0], pa[kk][1], pa[kk][2], pa[kk][3], b2, b3);
            }
        }
        __syncthreads();  // everyone done with this stage before it is refilled
    }
    cp_async_wait<0>();

    // ---- normalise and store
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        l_run[i] += __shfl_xor_sync(0xffffffffu, l_run[i], 1);
        l_run[i] += __shfl_xor_sync(0xffffffffu, l_run[i], 2);
    }
    const float inv0 = l_run[0] > 0.f ? 1.f / l_run[0] : 0.f;
    const float inv1 = l_run[1] > 0.f ? 1.f / l_run[1] : 0.f;
#pragma unroll
    for (int db = 0; db < D / 8; ++db) {
        const int col = db * 8 + 2 * tig;
        if (row_lo < n_q)
            *reinterpret_cast<uint32_t*>(out + (static_cast<long long>(q_begin + row_lo) * Hq + h) * D + col) =
                pack_bf16(o_acc[db][0] * inv0, o_acc[db][1] * inv0);
        if (row_hi < n_q)
            *reinterpret_cast<uint32_t*>(out + (static_cast<long long>(q_begin + row_hi) * Hq + h) * D + col) =
                pack_bf16(o_acc[db][2] * inv1, o_acc[db][3] * inv1);
    }
}

template <int D>
static int launch_prefill(const void* q, const void* kc, const void* vc, const int* bt, const int* cu_q,
                          const int* ctx, void* out, int S, int max_q_len, int Hq, int Hkv, int block_size,
                          int max_blocks, float scale, int flags, cudaStream_t stream) {
    constexpr int smem_bytes = kPfBM * D * 2 + 4 * kPfBN * D * 2;
    static bool attr_done[64] = {false};
    int dev = 0;
    ATX_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !attr_done[dev]) {
        ATX_CUDA(cudaFuncSetAttribute(attn_prefill_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
        attr_done[dev] = true;
    }
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = (flags & ATX_FLAG_PDL) ? 1 : 0;
    cfg.gridDim = dim3(static_cast<unsigned>((max_q_len + kPfBM - 1) / kPfBM), static_cast<unsigned>(Hq),
                       static_cast<unsigned>(S));
    cfg.blockDim = dim3(kPfThreads);
    cfg.dynamicSmemBytes = smem_bytes;
    cfg.stream = stream;
    ATX_CUDA(cudaLaunchKernelEx(&cfg, attn_prefill_kernel<D>, reinterpret_cast<const __nv_bfloat16*>(q),
                                reinterpret_cast<const __nv_bfloat16*>(kc), reinterpret_cast<const __nv_bfloat16*>(vc),
                                bt, cu_q, ctx, reinterpret_cast<__nv_bfloat16*>(out), Hq, Hkv, block_size, max_blocks,
                                scale * 1.4426950408889634f));
    return ATX_OK;
}

// attn_prefill_tc.cu
int launch_prefill_tc(const void* q, const void* kc, const void* vc, const int* bt, const int* cu_q, const int* ctx,
                      void* out, int S, int max_q_len, long long total_q, long long num_blocks, int Hq, int Hkv,
                      int max_blocks, float scale, int flags, cudaStream_t stream);

}  // namespace atx

using namespace atx;

extern "C" {

int atx_paged_attn_prefill(const void* q, const void* k_cache, const void* v_cache, const int* block_table,
                           const int* cu_q, const int* ctx_lens, void* out, int S, int max_q_len, long long total_q,
                           long long num_blocks, int Hq, int Hkv, int D, int block_size, int max_blocks, float scale,
                           int flags, atx_stream_t stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    ATX_REQUIRE(q && k_cache && v_cache && block_table && cu_q && ctx_lens && out,
                "atx_paged_attn_prefill: null pointer");
    ATX_REQUIRE(S > 0 && S <= 65535 && max_q_len > 0 && Hq > 0 && Hkv > 0 && Hq % Hkv == 0 && Hq <= 65535 &&
                    block_size > 0 && max_blocks > 0,
                "atx_paged_attn_prefill: bad shape S=%d max_q_len=%d Hq=%d Hkv=%d block_size=%d max_blocks=%d", S,
                max_q_len, Hq, Hkv, block_size, max_blocks);
    ATX_REQUIRE(total_q >= max_q_len && num_blocks > 0, "atx_paged_attn_prefill: total_q=%lld / num_blocks=%lld", total_q,
                num_blocks);
    if (D == 128 && block_size == 16 && (reinterpret_cast<uintptr_t>(q) & 15) == 0)   // the tcgen05 kernel
        return launch_prefill_tc(q, k_cache, v_cache, block_table, cu_q, ctx_lens, out, S, max_q_len, total_q, num_blocks,
                                 Hq, Hkv, max_blocks, scale, flags, stream);
    if (D == 128)
        return launch_prefill<128>(q, k_cache, v_cache, block_table, cu_q, ctx_lens, out, S, max_q_len, Hq, Hkv,
                                   block_size, max_blocks, scale, flags, stream);
    if (D == 64)
        return launch_prefill<64>(q, k_cache, v_cache, block_table, cu_q, ctx_lens, out, S, max_q_len, Hq, Hkv,
                                  block_size, max_blocks, scale, flags, stream);
    set_error("atx_paged_attn_prefill: unsupported head_dim=%d (64 or 128)", D);
    return ATX_ERR_UNSUPPORTED;
}

}  // extern "C"
