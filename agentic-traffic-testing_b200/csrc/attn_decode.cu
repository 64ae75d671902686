// Paged-KV decode attention (one query token per sequence), split along the KV length.
//
// Arithmetic follows eager_attention_forward (transformers/models/llama/modeling_llama.py:199-222):
// scores = q.k * head_dim**-0.5, softmax in fp32, out = P.V, with repeat_kv (:187-196) realised by
// letting the G = Hq/Hkv query heads of one KV head share every K/V row that is read.
//
// Decode attention is HBM-bound integer-indexed byte streaming: grid = (Hkv, B, splits); 16 lanes
// own one cache row (16-byte vector loads, rows of a page are contiguous so a warp reads whole
// 128-byte lines), score reductions are 4-step warp shuffles, the online softmax state lives in
// registers, and the next rows are already in flight (register double buffer) while the current
// ones are consumed. Partial (max, sum, out) per split go to caller scratch and are merged by
// attn_decode_reduce_kernel.
#include "common.cuh"

#include <cstdlib>
#include <mutex>

#include "../../include/atx_b200.h"

namespace atx {

constexpr int kDecThreads = 128;
constexpr int kDecWarps = kDecThreads / 32;
constexpr int kTokPerIter = kDecWarps * 2 * 2;  // warps x lane-groups x unroll

template <int DL>
struct RowVec;
template <>
struct RowVec<8> {
    uint4 v;
    __device__ __forceinline__ void load(const __nv_bfloat16* p) {
        asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
                     : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                     : "l"(p));
    }
    __device__ __forceinline__ void unpack(float (&f)[8]) const {
        f[0] = bf16lo(v.x); f[1] = bf16hi(v.x); f[2] = bf16lo(v.y); f[3] = bf16hi(v.y);
        f[4] = bf16lo(v.z); f[5] = bf16hi(v.z); f[6] = bf16lo(v.w); f[7] = bf16hi(v.w);
    }
};
template <>
struct RowVec<4> {
    uint2 v;
    __device__ __forceinline__ void load(const __nv_bfloat16* p) {
        asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
    }
    __device__ __forceinline__ void unpack(float (&f)[4]) const {
        f[0] = bf16lo(v.x); f[1] = bf16hi(v.x); f[2] = bf16lo(v.y); f[3] = bf16hi(v.y);
    }
};

__device__ __forceinline__ float group16_sum(float v) {
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    v += __shfl_xor_sync(0xffffffffu, v, 4);
    v += __shfl_xor_sync(0xffffffffu, v, 8);
    return v;
}
// exp2(a - b) with the convention that an empty state (a == -inf) contributes nothing
__device__ __forceinline__ float safe_exp2_diff(float a, float b) {
    return a == -INFINITY ? 0.f : exp2f(a - b);
}

template <int D, int G>
__global__ void __launch_bounds__(kDecThreads)
attn_decode_kernel(const __nv_bfloat16* __restrict__ q, const __nv_bfloat16* __restrict__ k_cache,
                   const __nv_bfloat16* __restrict__ v_cache, const int* __restrict__ block_table,
                   const int* __restrict__ seq_lens, __nv_bfloat16* __restrict__ out, float* __restrict__ ws_o,
                   float* __restrict__ ws_ml, int Hq, int Hkv, int block_size, int max_blocks, int num_splits,
                   float scale_log2) {
    constexpr int DL = D / 16;  // dims per lane
    __shared__ float sm_m[kDecWarps][G];
    __shared__ float sm_l[kDecWarps][G];
    __shared__ float sm_acc[kDecWarps][G][D];

    pdl_launch_dependents();
    pdl_wait();
    const int g = blockIdx.x, b = blockIdx.y, sp = blockIdx.z;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int grp = lane >> 4, sub = lane & 15;
    const int L = seq_lens[b];
    int chunk = (L + num_splits - 1) / num_splits;
    chunk = (chunk + kTokPerIter - 1) / kTokPerIter * kTokPerIter;
    const int begin = sp * chunk;
    const int end = min(L, begin + chunk);

    float qf[G][DL];
#pragma unroll
    for (int j = 0; j < G; ++j) {
        RowVec<DL> r;
        r.load(q + (static_cast<long long>(b) * Hq + g * G + j) * D + sub * DL);
        r.unpack(qf[j]);
#pragma unroll
        for (int d = 0; d < DL; ++d) qf[j][d] *= scale_log2;
    }
    float m[G], l[G], acc[G][DL];
#pragma unroll
    for (int j = 0; j < G; ++j) {
        m[j] = -INFINITY;
        l[j] = 0.f;
#pragma unroll
        for (int d = 0; d < DL; ++d) acc[j][d] = 0.f;
    }

    const int* bt = block_table + static_cast<long long>(b) * max_blocks;
    const long long head_stride = static_cast<long long>(block_size) * D;  // elements per (block, head)
    auto row_offset = [&](int tok) -> long long {
        const int c = min(tok, end - 1);  // clamp: masked rows still read a valid address
        const int page = __ldg(bt + c / block_size);
        return (static_cast<long long>(page) * Hkv + g) * head_stride + static_cast<long long>(c % block_size) * D +
               sub * DL;
    };

    // Loop bounds are warp-uniform (the shuffles below need all 32 lanes); rows past `end` are
    // read from a clamped, valid address and masked to -inf.
    int wbase = begin + warp * 4;
    RowVec<DL> k0, k1, v0, v1;
    if (wbase < end) {
        const long long o0 = row_offset(wbase + grp * 2), o1 = row_offset(wbase + grp * 2 + 1);
        k0.load(k_cache + o0);
        k1.load(k_cache + o1);
        v0.load(v_cache + o0);
        v1.load(v_cache + o1);
    }
    for (; wbase < end; wbase += kTokPerIter) {
        const int base = wbase + grp * 2;
        RowVec<DL> nk0, nk1, nv0, nv1;
        if (wbase + kTokPerIter < end) {
            const long long o0 = row_offset(base + kTokPerIter), o1 = row_offset(base + kTokPerIter + 1);
            nk0.load(k_cache + o0);
            nk1.load(k_cache + o1);
            nv0.load(v_cache + o0);
            nv1.load(v_cache + o1);
        }
        const bool valid0 = base < end, valid1 = base + 1 < end;
        float kf0[DL], kf1[DL];
        k0.unpack(kf0);
        k1.unpack(kf1);
        float s0[G], s1[G];
#pragma unroll
        for (int j = 0; j < G; ++j) {
            float a = 0.f, c = 0.f;
#pragma unroll
            for (int d = 0; d < DL; ++d) {
                a = fmaf(qf[j][d], kf0[d], a);
                c = fmaf(qf[j][d], kf1[d], c);
            }
            a = group16_sum(a);
            c = group16_sum(c);
            s0[j] = valid0 ? a : -INFINITY;
            s1[j] = valid1 ? c : -INFINITY;
        }
        float vf0[DL], vf1[DL];
        v0.unpack(vf0);
        v1.unpack(vf1);
#pragma unroll
        for (int j = 0; j < G; ++j) {
            const float mn = fmaxf(m[j], fmaxf(s0[j], s1[j]));
            const float mref = mn == -INFINITY ? 0.f : mn;  // nothing seen yet: every factor below is 0
            const float corr = exp2f(m[j] - mref);
            const float p0 = exp2f(s0[j] - mref);
            const float p1 = exp2f(s1[j] - mref);
            m[j] = mn;
            l[j] = l[j] * corr + p0 + p1;
#pragma unroll
            for (int d = 0; d < DL; ++d) acc[j][d] = fmaf(p1, vf1[d], fmaf(p0, vf0[d], acc[j][d] * corr));
        }
        k0 = nk0;
        k1 = nk1;
        v0 = nv0;
        v1 = nv1;
    }

    // merge the two lane groups of the warp
#pragma unroll
    for (int j = 0; j < G; ++j) {
        const float mo = __shfl_xor_sync(0xffffffffu, m[j], 16);
        const float lo = __shfl_xor_sync(0xffffffffu, l[j], 16);
        const float mn = fmaxf(m[j], mo);
        const float fs = safe_exp2_diff(m[j], mn), fo = safe_exp2_diff(mo, mn);
        l[j] = l[j] * fs + lo * fo;
#pragma unroll
        for (int d = 0; d < DL; ++d) {
            const float ao = __shfl_xor_sync(0xffffffffu, acc[j][d], 16);
            acc[j][d] = acc[j][d] * fs + ao * fo;
        }
        m[j] = mn;
    }
    if (grp == 0) {
#pragma unroll
        for (int j = 0; j < G; ++j) {
            if (sub == 0) {
                sm_m[warp][j] = m[j];
                sm_l[warp][j] = l[j];
            }
#pragma unroll
            for (int d = 0; d < DL; ++d) sm_acc[warp][j][sub * DL + d] = acc[j][d];
        }
    }
    __syncthreads();
    // merge the warps; one (head, dim) item per thread-iteration
    for (int it = threadIdx.x; it < G * D; it += kDecThreads) {
        const int j = it / D, d = it % D;
        float M = sm_m[0][j];
#pragma unroll
        for (int w = 1; w < kDecWarps; ++w) M = fmaxf(M, sm_m[w][j]);
        float o = 0.f, ls = 0.f;
#pragma unroll
        for (int w = 0; w < kDecWarps; ++w) {
            const float f = safe_exp2_diff(sm_m[w][j], M);
            o += sm_acc[w][j][d] * f;
            ls += sm_l[w][j] * f;
        }
        const long long h = static_cast<long long>(b) * Hq + g * G + j;
        if (num_splits == 1) {
            out[h * D + d] = __float2bfloat16_rn(ls > 0.f ? o / ls : 0.f);
        } else {
            ws_o[(h * num_splits + sp) * D + d] = o;
            if (d == 0) {
                ws_ml[(h * num_splits + sp) * 2] = M;
                ws_ml[(h * num_splits + sp) * 2 + 1] = ls;
            }
        }
    }
}

// ---- tensor-core decode attention --------------------------------------------------------------------
// The CUDA-core kernel above issues ~90 warp instructions per (token, kv-head) row of 512 bytes; at
// 6.5 TB/s that is the whole issue bandwidth of the chip (ncu: issue slots 65 % busy, DRAM 7 %), so it
// can never be HBM-bound. Here the G query heads of one KV head become rows of a 16-row MMA tile
// (rows >= G are zero padding): QK^T and PV run on mma.sync m16n8k16 and one warp needs ~8 instructions
// per row. One WARP per CTA (no block-level merge), K/V pages staged with cp.async into a 3-deep ring
// of XOR-swizzled 16-token tiles, read back with ldmatrix. Requires block_size % 16 == 0 so that a
// tile never straddles a page; other block sizes keep using attn_decode_kernel.
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
    // Sequence lengths, block tables and every K/V row except the one written by this step's
    // rope_kv_write (always in the last tile of the sequence) do not depend on the upstream kernel:
    // those tiles are requested BEFORE griddepcontrol.wait. Commit groups stay in tile order.
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

// ---- cluster variant: the KV splits of one (sequence, KV head) form a thread-block CLUSTER ------------
// Short-context decode attention is a latency chain, not a bandwidth problem (in-graph ablation: 9 us
// per layer against 4 us of HBM time), so this variant removes links from the chain:
//   * everything that does not depend on the upstream kernel is issued BEFORE griddepcontrol.wait:
//     sequence length, block-table lookups and the cp.async of every K/V tile except the one that
//     holds the token written by this step's rope_kv_write (always the last tile of the sequence);
//   * the per-split partials never touch global memory and there is no second kernel: each CTA parks
//     (m, l, o) in its shared memory, one barrier.cluster, then CTA r merges slice r of the output by
//     reading its peers through distributed shared memory (mapa + ld.shared::cluster).
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ float dsmem_ld_f32(const float* local_ptr, uint32_t cta) {
    uint32_t remote;
    float v;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(local_ptr)), "r"(cta));
    asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(remote) : "memory");
    return v;
}

constexpr int kClStages = 4;  // 3 of the ~5 tiles of a 72-token split are in flight before the wait

template <int D>
__global__ void __launch_bounds__(32)
attn_decode_cluster_kernel(const __nv_bfloat16* __restrict__ q, const __nv_bfloat16* __restrict__ k_cache,
                           const __nv_bfloat16* __restrict__ v_cache, const int* __restrict__ block_table,
                           const int* __restrict__ seq_lens, __nv_bfloat16* __restrict__ out, int Hq, int Hkv,
                           int block_size, int max_blocks, int num_splits, float scale_log2, int early_prefetch) {
    constexpr int CH = D / 8;
    constexpr int TILE_BYTES = kMmaTile * D * 2;
    constexpr int kRedFloats = 8 * (D + 2);  // up to 8 heads: o[G][D], m[G], l[G]
    __shared__ __align__(128) uint8_t smem[kClStages * 2 * TILE_BYTES];
    __shared__ float red[kRedFloats];

    pdl_launch_dependents();
    if (!early_prefetch) pdl_wait();
    const int g = blockIdx.x, b = blockIdx.y, sp = blockIdx.z;
    const int lane = threadIdx.x;
    const int gid = lane >> 2, tig = lane & 3;
    const int G = Hq / Hkv;
    // host-staged inputs: independent of the upstream kernel
    const int L = seq_lens[b];
    int chunk = (L + num_splits - 1) / num_splits;
    chunk = (chunk + kMmaTile - 1) / kMmaTile * kMmaTile;
    const int begin = sp * chunk;
    const int end = min(L, begin + chunk);
    const int n_tiles = end > begin ? (end - begin + kMmaTile - 1) / kMmaTile : 0;
    const int* bt = block_table + static_cast<long long>(b) * max_blocks;
    // the tile that holds token L-1 (written by this step) may only be read after the wait
    const int fresh_tile = (end == L && n_tiles > 0) ? n_tiles - 1 : -1;

    auto issue_tile = [&](int t) {
        const int tok0 = begin + t * kMmaTile;
        const int page = __ldg(bt + tok0 / block_size);
        const long long base = ((static_cast<long long>(page) * Hkv + g) * block_size + tok0 % block_size) * D;
        uint8_t* sk = smem + (t % kClStages) * 2 * TILE_BYTES;
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
    // prologue: commit groups stay in tile order. Old tiles go out before the wait; the fresh tile (always
    // the last tile of the sequence) and the empty groups that pad the ring follow after it.
    int t_pre = 0;
    for (; t_pre < kClStages - 1; ++t_pre) {
        if (t_pre >= n_tiles || t_pre == fresh_tile) break;
        issue_tile(t_pre);
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    pdl_wait();
    for (int t = t_pre; t < kClStages - 1; ++t) {
        if (t < n_tiles) issue_tile(t);
        asm volatile("cp.async.commit_group;" ::: "memory");
    }

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
        if (t + kClStages - 1 < n_tiles) issue_tile(t + kClStages - 1);
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group %0;" ::"n"(kClStages - 1) : "memory");
        __syncwarp();
        const uint8_t* sk = smem + (t % kClStages) * 2 * TILE_BYTES;
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
        const float mn = fmaxf(m_run, mx);
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
        __syncwarp();
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");

    // ---- park this split's (o, m, l) in shared memory, then merge across the cluster through DSMEM
    l_run += __shfl_xor_sync(0xffffffffu, l_run, 1);
    l_run += __shfl_xor_sync(0xffffffffu, l_run, 2);
    if (gid < G) {
#pragma unroll
        for (int db = 0; db < D / 8; ++db) {
            red[gid * D + db * 8 + 2 * tig] = o_acc[db][0];
            red[gid * D + db * 8 + 2 * tig + 1] = o_acc[db][1];
        }
        if (tig == 0) {
            red[G * D + gid] = m_run == -INFINITY ? -INFINITY : m_run * scale_log2;
            red[G * D + G + gid] = l_run;
        }
    }
    __syncwarp();
    cluster_sync_all();
    const int S = num_splits;
    const int E = G * D;                       // output elements of this (sequence, KV head)
    const int per = (E + S - 1) / S;           // slice of CTA `sp`
    for (int e = sp * per + lane; e < min(E, (sp + 1) * per); e += 32) {
        const int j = e / D;
        float M = -INFINITY;
        for (int s2 = 0; s2 < S; ++s2) M = fmaxf(M, dsmem_ld_f32(&red[G * D + j], s2));
        float o = 0.f, ls = 0.f;
        for (int s2 = 0; s2 < S; ++s2) {
            const float f = safe_exp2_diff(dsmem_ld_f32(&red[G * D + j], s2), M);
            o += dsmem_ld_f32(&red[e], s2) * f;
            ls += dsmem_ld_f32(&red[G * D + G + j], s2) * f;
        }
        const long long h = static_cast<long long>(b) * Hq + g * G + j;
        out[h * D + (e - j * D)] = __float2bfloat16_rn(ls > 0.f ? o / ls : 0.f);
    }
    __syncwarp();
    cluster_sync_all();  // nobody leaves while a peer may still be reading its shared memory
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

// The cluster variant is opt-in (ATX_DECODE_CLUSTER=1). Measured on B200 (r01): faster in isolation
// (14.8 vs 18.9 us at 6 x 1152 tokens) but slower inside the decode graph (3.07 vs 2.91 ms per step):
// 16-CTA clusters have to be co-scheduled on one GPC while PDL-launched GEMM CTAs hold the shared memory.
static const bool g_decode_cluster = [] {
    const char* e = getenv("ATX_DECODE_CLUSTER");
    return e && e[0] == '1';
}();

static const bool g_decode_early_prefetch = [] {
    const char* e = getenv("ATX_DECODE_PREFETCH");
    return !(e && e[0] == '0');
}();

template <int D, int G>
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
    cfg.stream = stream;
    if (block_size % kMmaTile == 0 && G <= 8 && num_splits > 1 && num_splits <= 16 && g_decode_cluster) {
        static bool attr_done[64] = {false};
        int dev = 0;
        ATX_CUDA(cudaGetDevice(&dev));
        if (dev >= 0 && dev < 64 && !attr_done[dev]) {
            ATX_CUDA(cudaFuncSetAttribute(attn_decode_cluster_kernel<D>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
            attr_done[dev] = true;
        }
        cudaLaunchAttribute cattr[2];
        cattr[0].id = cudaLaunchAttributeClusterDimension;
        cattr[0].val.clusterDim.x = 1;
        cattr[0].val.clusterDim.y = 1;
        cattr[0].val.clusterDim.z = static_cast<unsigned>(num_splits);
        cattr[1] = attr[0];
        cfg.attrs = cattr;
        cfg.numAttrs = (flags & ATX_FLAG_PDL) ? 2 : 1;
        cfg.blockDim = dim3(32);
        ATX_CUDA(cudaLaunchKernelEx(&cfg, attn_decode_cluster_kernel<D>, reinterpret_cast<const __nv_bfloat16*>(q),
                                    reinterpret_cast<const __nv_bfloat16*>(kc),
                                    reinterpret_cast<const __nv_bfloat16*>(vc), bt, sl,
                                    reinterpret_cast<__nv_bfloat16*>(out), Hq, Hkv, block_size, max_blocks, num_splits,
                                    scale * 1.4426950408889634f, g_decode_early_prefetch ? 1 : 0));
        return ATX_OK;
    }
    if (block_size % kMmaTile == 0 && G <= 8) {
        cfg.blockDim = dim3(32);
        ATX_CUDA(cudaLaunchKernelEx(&cfg, attn_decode_mma_kernel<D>, reinterpret_cast<const __nv_bfloat16*>(q),
                                    reinterpret_cast<const __nv_bfloat16*>(kc),
                                    reinterpret_cast<const __nv_bfloat16*>(vc), bt, sl,
                                    reinterpret_cast<__nv_bfloat16*>(out), ws_o, ws_ml, Hq, Hkv, block_size, max_blocks,
                                    num_splits, scale * 1.4426950408889634f));
    } else {
    cfg.blockDim = dim3(kDecThreads);
    ATX_CUDA(cudaLaunchKernelEx(&cfg, attn_decode_kernel<D, G>, reinterpret_cast<const __nv_bfloat16*>(q),
                                reinterpret_cast<const __nv_bfloat16*>(kc), reinterpret_cast<const __nv_bfloat16*>(vc),
                                bt, sl, reinterpret_cast<__nv_bfloat16*>(out), ws_o, ws_ml, Hq, Hkv, block_size,
                                max_blocks, num_splits, scale * 1.4426950408889634f));
    }
    if (num_splits > 1) {
        cfg.gridDim = dim3(static_cast<unsigned>(B * Hq));
        cfg.blockDim = dim3(static_cast<unsigned>(D));
        ATX_CUDA(cudaLaunchKernelEx(&cfg, attn_decode_reduce_kernel, static_cast<const float*>(ws_o),
                                    static_cast<const float*>(ws_ml), reinterpret_cast<__nv_bfloat16*>(out), D,
                                    num_splits));
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
    ATX_REQUIRE(B > 0 && Hq > 0 && Hkv > 0 && Hq % Hkv == 0 && block_size > 0 && max_blocks > 0 && num_splits > 0,
                "atx_paged_attn_decode: bad shape B=%d Hq=%d Hkv=%d block_size=%d max_blocks=%d splits=%d", B, Hq,
                Hkv, block_size, max_blocks, num_splits);
    ATX_REQUIRE(num_splits <= 65535 && B <= 65535, "atx_paged_attn_decode: B and num_splits must be <= 65535");
    float* ws_o = nullptr;
    float* ws_ml = nullptr;
    if (num_splits > 1) {
        const long long need = atx_paged_attn_decode_workspace_bytes(B, Hq, D, num_splits);
        ATX_REQUIRE(workspace && workspace_bytes >= need,
                    "atx_paged_attn_decode: workspace of %lld bytes needed, %lld given", need, workspace_bytes);
        ws_o = static_cast<float*>(workspace);
        ws_ml = ws_o + static_cast<long long>(B) * Hq * num_splits * D;
    }
    const int G = Hq / Hkv;
#define ATX_DEC_CASE(DD, GG)                                                                                   \
    if (D == DD && G == GG)                                                                                    \
        return launch_decode<DD, GG>(q, kc, vc, bt, sl, out, ws_o, ws_ml, B, Hq, Hkv, block_size, max_blocks, \
                                     num_splits, scale, flags, stream);
    ATX_DEC_CASE(128, 4)
    ATX_DEC_CASE(128, 8)
    ATX_DEC_CASE(64, 4)
    ATX_DEC_CASE(128, 1)
    ATX_DEC_CASE(128, 2)
    ATX_DEC_CASE(64, 1)
    ATX_DEC_CASE(64, 2)
    ATX_DEC_CASE(64, 8)
#undef ATX_DEC_CASE
    set_error("atx_paged_attn_decode: unsupported head_dim=%d / group=%d (head_dim 64|128, group 1|2|4|8)", D, G);
    return ATX_ERR_UNSUPPORTED;
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
    ATX_REQUIRE(B > 0 && Hq > 0 && Hkv > 0 && Hq % Hkv == 0 && block_size > 0 && max_blocks > 0 && num_splits > 0,
                "atx_paged_attn_decode_static: bad shape B=%d Hq=%d Hkv=%d block_size=%d max_blocks=%d splits=%d", B,
                Hq, Hkv, block_size, max_blocks, num_splits);
    ATX_REQUIRE(num_splits <= 65535 && B <= 65535, "atx_paged_attn_decode_static: B and num_splits must be <= 65535");
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
    const int G = Hq / Hkv;
#define ATX_DEC_STATIC_CASE(DD, GG)                                                                           \
    if (D == DD && G == GG)                                                                                    \
        return launch_decode<DD, GG>(q, kc, vc, bt, sl, out, ws_o, ws_ml, B, Hq, Hkv, block_size, max_blocks, \
                                     num_splits, scale, flags, stream);
    ATX_DEC_STATIC_CASE(128, 4)
    ATX_DEC_STATIC_CASE(128, 8)
    ATX_DEC_STATIC_CASE(64, 4)
    ATX_DEC_STATIC_CASE(128, 1)
    ATX_DEC_STATIC_CASE(128, 2)
    ATX_DEC_STATIC_CASE(64, 1)
    ATX_DEC_STATIC_CASE(64, 2)
    ATX_DEC_STATIC_CASE(64, 8)
#undef ATX_DEC_STATIC_CASE
    set_error("atx_paged_attn_decode_static: unsupported head_dim=%d / group=%d (head_dim 64|128, group 1|2|4|8)", D,
              G);
    return ATX_ERR_UNSUPPORTED;
}

}  // extern "C"
