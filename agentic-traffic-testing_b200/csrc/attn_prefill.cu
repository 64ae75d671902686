// Causal prefill attention over the paged KV cache on tcgen05 tensor cores (head_dim 64 | 128, 16-token pages).
//
// Arithmetic follows eager_attention_forward (transformers/models/llama/modeling_llama.py:199-222) with the additive
// causal mask, GQA by head sharing (repeat_kv :187-196) and fp32 softmax (:217). The chunk's own K/V are already in the
// pages (atx_rope_kv_write ran first), so every key is read through the block table: prefix-cache hits and chunked
// prefill need no special case. Both GEMMs of attention run on tcgen05 with TMEM accumulators (head_dim 128; 64 for the
// 1B model: one 64-dim atom per tile instead of two, N = 64 for P.V):
//
//   S  = Q K^T   M=128 query rows, N=64 keys,  K=128 dims     A=Q (smem, K-major)  B=K  (smem, K-major)
//   O += P V     M=128 query rows, N=128 dims, K=64 keys      A=P (TMEM)           B=V  (smem, MN-major)
//
// One CTA = TWO 128-row query tiles (A: rows m0.., B: rows m0+128..) of one head; K/V arrive in 128-key tiles and are
// handed between the tensor pipe and the softmax warps in 64-key HALVES, which gives FOUR independent chains
// (tile A | B) x (keys 0-63 | 64-127):   [wait P_c(j)]  P.V_c(j)  QK_c(j+1)  ->  S_c(j+1)  ->  softmax  ->  P_c(j+1)
// While the softmax warps of a tile work on one half, the tensor pipe runs the other half and the other tile. Roles:
//   warp 0      TMA producer: Q_A, Q_B once; K and V in 128-key tiles (eight pages; a page of one KV head is
//               16 rows x 256 B, two 128B-swizzled boxes, lane p loads page p), 2-deep rings, K one tile ahead of V
//   warp 1      the single MMA-issuing thread, chains in the fixed rotation A_lo, B_lo, A_hi, B_hi. QK_c(j+1)
//               overwrites exactly the TMEM columns P_c(j) sits in: tcgen05.mma execute in issue order, so that is
//               safe once P.V_c(j) has been issued, and S needs no second buffer
//   warps 2-5   softmax of tile A, warps 6-9 of tile B: ONE thread per query row (no cross-thread exchange):
//               64 scores from TMEM in one round trip, causal mask on diagonal tiles, running max with a lazy
//               reference (O in TMEM is rescaled only when some row of the warp grew by more than 2^8; then the
//               P.V of the other half, possibly still accumulating, is waited for first), ex2 with packed
//               FFMA2 / FADD2 arithmetic around it, P as bf16 pairs written back into the first 32 columns of the
//               half just read (tcgen05.st): P.V takes its A operand from TMEM, P never crosses shared memory;
//               finally O / l -> bf16
// TMEM (512 columns): S_A, S_B (128 each = two 64-column halves), O_A, O_B (128 each). K and V tiles are shared by
// both query tiles. Grid: 1-D, the G heads of a KV group fastest, then the query tile pairs of a (sequence, KV head)
// from the LAST (most keys) to the first: the launch ends on short CTAs and the resident CTAs share K/V in L2.
// History (profiles/r01n, r02j, r02zc): one tile per CTA with two threads per row: tensor pipe 31 % active; two
// tiles with 64-key tiles and double-buffered S: 47.5 % (each P store waited for the previous P.V of its tile: a
// ~1500-cycle serial chain per 64 keys); 128-key tiles, two chains: 57 % (softmax of a tile 2200 cycles against a
// 1100-cycle MMA batch); this version: 61 %, 1205 TFLOP/s causal at S=4 x 8192. What is left: the MUFU unit
// retires 4 ex2 per cycle and scheduler (XU pipe 62 % busy), the softmax warps wait 34 % of their time for S, and a
// pure MMA + TMA stream of this operand mix without any hand-over reaches 1555 TFLOP/s at the 1.64 GHz the SMs hold.
// Tried and dropped here: a second MMA-issuing thread (1121), a polynomial exp2 for a share of the scores (slower at
// every share: it costs more issue slots than the MUFU it frees), 128 threads arriving vs one per warp (same).
#include "common.cuh"

#include <mutex>
#include <vector>

#include "../../include/atx_b200.h"

namespace atx {

constexpr int kTcBM = 128;        // query rows per tile (two tiles per CTA)
constexpr int kTcBN = 128;        // keys per K/V tile
constexpr int kTcHalf = 64;       // keys per softmax / MMA hand-over (half a tile)
constexpr int kTcThreads = 320;   // TMA warp, MMA warp, 4 + 4 softmax warps
constexpr int kTcPage = 16;
constexpr int kQAtom = 128 * 128;               // [128 rows x 64 bf16], 128B-swizzled: one 64-dim atom of Q
constexpr int kKVAtom = kTcBN * 128;            // [128 keys x 64 bf16]: one 64-dim atom of a K or V tile
constexpr int kKS = 2, kVS = 2;                 // ring depths
template <int D> struct TcCfg {                 // head_dim 64 (one dim-atom per tile) or 128 (two)
    static constexpr int kAtoms = D / 64;
    static constexpr int kQTile = kAtoms * kQAtom;      // [128 rows x D] bf16
    static constexpr int kKVTile = kAtoms * kKVAtom;    // [128 keys x D] bf16
    static constexpr int kSmem = 2 * kQTile + (kKS + kVS) * kKVTile + 1024 + 512;   // + align + barriers
};

// MN-major operand ([K rows][64 contiguous N elements] atoms, 128B swizzle): LBO = distance between the
// N-atoms, SBO = distance between groups of 8 K-rows (cute::UMMA canonical Major-MN SW128 layout).
__device__ __forceinline__ uint64_t umma_desc_sw128_mn(uint32_t smem_addr, uint32_t lbo_bytes) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
    d |= static_cast<uint64_t>(lbo_bytes >> 4) << 16;
    d |= static_cast<uint64_t>(1024 >> 4) << 32;
    d |= static_cast<uint64_t>(1) << 46;
    d |= static_cast<uint64_t>(2) << 61;
    return d;
}
__device__ __forceinline__ float ex2_approx(float x) {  // one MUFU op; 2^-inf = 0
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// Packed fp32 pairs (FFMA2 / FADD2 on sm_100): the softmax warps are issue-bound, two scores per instruction.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
    f32x2 v;
    asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(lo), "f"(hi));
    return v;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

template <int kTcD>
__global__ void __launch_bounds__(kTcThreads, 1)
attn_prefill_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                       const __grid_constant__ CUtensorMap tmV, const int* __restrict__ block_table,
                       const int* __restrict__ cu_q, const int* __restrict__ ctx_lens,
                       __nv_bfloat16* __restrict__ out, int Hq, int Hkv, int max_blocks, int n_mtiles,
                       float scale_log2) {
    constexpr int kAtoms = TcCfg<kTcD>::kAtoms, kQTile = TcCfg<kTcD>::kQTile, kKVTile = TcCfg<kTcD>::kKVTile;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    uint8_t* sQ = smem;                                   // tile A, tile B
    uint8_t* sK = sQ + 2 * kQTile;                        // kKS stages
    uint8_t* sV = sK + kKS * kKVTile;                     // kVS stages
    uint64_t* bars = reinterpret_cast<uint64_t*>(sV + kVS * kKVTile);
    uint64_t* q_full = bars;            // 1
    uint64_t* k_full = bars + 1;        // 4 slots reserved each
    uint64_t* k_empty = bars + 5;
    uint64_t* v_full = bars + 9;
    uint64_t* v_empty = bars + 13;
    uint64_t* s_full = bars + 17;       // [tile][half]: 4
    uint64_t* p_full = bars + 21;       // [tile][half]: 4
    uint64_t* o_done = bars + 25;       // [tile][half]: 4
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 29);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // 1-D grid, fastest to slowest: the G query heads of one KV head | query tile pair, LAST first | KV head | sequence.
    // Causal: later query tiles see more keys, so each (sequence, KV head) is worked heaviest tile first and the tail of
    // the launch is made of short CTAs; the CTAs resident at one time share one or two KV heads' keys (L2 hits).
    const int G = Hq / Hkv;
    const int bx = blockIdx.x;
    const int kvs = bx / (G * n_mtiles);
    const int s = kvs / Hkv, h = (kvs % Hkv) * G + bx % G;
    const int q_begin = cu_q[s];
    const int n_q = cu_q[s + 1] - q_begin;
    const int m0 = (n_mtiles - 1 - (bx / G) % n_mtiles) * 2 * kTcBM;
    if (m0 >= n_q) return;  // whole CTA: before any barrier / TMEM use
    const int ctx = ctx_lens[s];
    const int off = ctx - n_q;  // query j sees keys [0, off + j]
    const int g = h / G;
    const bool has_b = m0 + kTcBM < n_q;                      // the second tile has at least one real row
    // keys each tile needs, in 64-key tiles (B's rows see up to 128 more keys than A's)
    const int nt_a = (min(ctx, off + m0 + kTcBM) + kTcBN - 1) / kTcBN;
    const int nt_b = has_b ? (min(ctx, off + m0 + 2 * kTcBM) + kTcBN - 1) / kTcBN : 0;
    const int n_tiles = max(nt_a, nt_b);
    const int* bt = block_table + static_cast<long long>(s) * max_blocks;

    if (warp == 0 && elect_one()) {
        tma_prefetch_desc(&tmQ);
        tma_prefetch_desc(&tmK);
        tma_prefetch_desc(&tmV);
        mbar_init(q_full, 1);
        for (int i = 0; i < 4; ++i) {
            mbar_init(&k_full[i], 1);
            mbar_init(&k_empty[i], 1);
            mbar_init(&v_full[i], 1);
            mbar_init(&v_empty[i], 1);
            mbar_init(&s_full[i], 1);
            mbar_init(&p_full[i], 128);
            mbar_init(&o_done[i], 1);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<512>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // columns: S_A 0, S_B 128, each as two 64-key halves of 64 columns (P of a half = bf16 pairs in the first 32
    // columns of that half), O_A 256, O_B 384
    pdl_launch_dependents();

    if (warp == 0) {
        // ===== TMA producer: the whole warp works. Lane p < 8 owns page p of a tile: the block-table lookups of
        // a tile are one coalesced load, two tiles ahead of their use, and lane p issues its page's boxes.
        pdl_wait();  // q and this chunk's K/V rows come from the upstream kernels
        if (lane == 0) {
            mbar_expect_tx(q_full, has_b ? 2 * kQTile : kQTile);
            for (int t = 0; t < (has_b ? 2 : 1); ++t) {
#pragma unroll
                for (int a = 0; a < kAtoms; ++a)
                    tma_load_2d(sQ + t * kQTile + a * kQAtom, &tmQ, q_full, h * kTcD + a * 64, q_begin + m0 + t * kTcBM,
                                kEvictNormal);
            }
        }
        auto tile_pages = [&](int j) { return min(kTcBN / kTcPage, (ctx - j * kTcBN + kTcPage - 1) / kTcPage); };
        auto page_row = [&](int j) -> int {   // first cache row of this lane's page in tile j (or -1)
            return (j < n_tiles && lane < tile_pages(j))
                       ? (__ldg(bt + j * (kTcBN / kTcPage) + lane) * Hkv + g) * kTcPage : -1;
        };
        auto load = [&](int j, int row, uint8_t* ring, int depth, uint64_t* full, uint64_t* empty,
                        const CUtensorMap* tm) {
            const int st = j % depth;
            if (lane == 0) {
                mbar_wait(&empty[st], ((j / depth) & 1) ^ 1);
                mbar_expect_tx(&full[st], tile_pages(j) * kTcPage * kTcD * 2);   // allocated pages only
            }
            __syncwarp();
            if (row >= 0) {
                uint8_t* dst = ring + st * kKVTile + lane * kTcPage * 128;
#pragma unroll
                for (int a = 0; a < kAtoms; ++a) tma_load_2d(dst + a * kKVAtom, tm, &full[st], a * 64, row, kEvictLast);
            }
        };
        // K runs one tile ahead of V (QK^T of tile j+1 is issued before P.V of tile j): K0, K1, V0, K2, V1, ...
        int row_cur = page_row(0);
        int row_nxt = page_row(1);
        load(0, row_cur, sK, kKS, k_full, k_empty, &tmK);
        for (int j = 0; j < n_tiles; ++j) {
            const int row_nn = page_row(j + 2);
            if (j + 1 < n_tiles) load(j + 1, row_nxt, sK, kKS, k_full, k_empty, &tmK);
            load(j, row_cur, sV, kVS, v_full, v_empty, &tmV);
            row_cur = row_nxt;
            row_nxt = row_nn;
        }
    } else if (warp == 1) {
        if (elect_one()) {
            // ===== MMA issuer. Four independent chains in rotation, (tile A | B) x (keys 0-63 | 64-127 of the 128-key
            // K/V tile): for chain c and tile j   [wait P_c(j)]  P.V_c(j)  QK_c(j+1)  -> S_c(j+1)
            // QK_c(j+1) overwrites exactly the columns P_c(j) sits in; tcgen05.mma execute in issue order, so that is
            // safe once P.V_c(j) has been issued. While the softmax warps of a tile work on one half, the tensor pipe
            // runs the three other chains: neither side waits for a whole tile of the other.
            constexpr uint32_t idesc_qk = umma_idesc_bf16(kTcBM, kTcHalf);               // A, B K-major
            constexpr uint32_t idesc_pv = umma_idesc_bf16(kTcBM, kTcD) | (1u << 16);      // B (= V) MN-major
            auto issue_qk = [&](int t, int hf, int j) {
                const uint32_t ka = smem_u32(sK + (j % kKS) * kKVTile) + hf * kTcHalf * 128;
                const uint32_t qa = smem_u32(sQ + t * kQTile);
#pragma unroll
                for (int kk = 0; kk < kTcD / 16; ++kk)
                    umma_bf16(tmem_base + t * 128 + hf * kTcHalf,
                              umma_desc_sw128(qa + (kk >> 2) * kQAtom + (kk & 3) * 32),
                              umma_desc_sw128(ka + (kk >> 2) * kKVAtom + (kk & 3) * 32), idesc_qk, kk > 0 ? 1u : 0u);
                umma_commit(&s_full[t * 2 + hf]);
            };
            mbar_wait(q_full, 0);
            mbar_wait(&k_full[0], 0);
            tc_fence_after();
#pragma unroll
            for (int hf = 0; hf < 2; ++hf) {
                if (0 < nt_a) issue_qk(0, hf, 0);
                if (0 < nt_b) issue_qk(1, hf, 0);
            }
            umma_commit(&k_empty[0]);
            for (int j = 0; j < n_tiles; ++j) {
                const int vst = j % kVS;
                bool v_ready = false, k_ready = false;
#pragma unroll
                for (int hf = 0; hf < 2; ++hf) {
#pragma unroll
                    for (int t = 0; t < 2; ++t) {
                        const int nt = t == 0 ? nt_a : nt_b;
                        if (j < nt) {
                            if (!v_ready) {
                                mbar_wait(&v_full[vst], (j / kVS) & 1);
                                v_ready = true;
                            }
                            mbar_wait(&p_full[t * 2 + hf], j & 1);   // P in TMEM, O rescaled, tail rows of V cleared
                            tc_fence_after();
                            const uint32_t va = smem_u32(sV + vst * kKVTile) + hf * kTcHalf * 128;
#pragma unroll
                            for (int kk = 0; kk < kTcHalf / 16; ++kk)
                                umma_bf16_ts(tmem_base + 256 + t * 128, tmem_base + t * 128 + hf * kTcHalf + kk * 8,
                                             umma_desc_sw128_mn(va + kk * 16 * 128, kKVAtom), idesc_pv,
                                             (j > 0 || hf > 0 || kk > 0) ? 1u : 0u);
                            umma_commit(&o_done[t * 2 + hf]);
                        }
                        if (j + 1 < nt) {
                            if (!k_ready) {
                                mbar_wait(&k_full[(j + 1) % kKS], ((j + 1) / kKS) & 1);
                                tc_fence_after();
                                k_ready = true;
                            }
                            issue_qk(t, hf, j + 1);
                        }
                    }
                }
                umma_commit(&v_empty[vst]);
                if (j + 1 < n_tiles) umma_commit(&k_empty[(j + 1) % kKS]);
            }
        }
    } else {
        // ===== softmax / correction / epilogue: warps 2-5 own tile A, warps 6-9 tile B; a warp reads the TMEM
        // lane quarter (warp & 3), thread = one query row, all 64 keys of the tile in its registers.
        const int t = (warp - 2) >> 2;                   // 0 = tile A, 1 = tile B
        const int nt = t == 0 ? nt_a : nt_b;
        const int quarter = warp & 3;
        const int row = quarter * 32 + lane;             // row inside the tile == TMEM lane
        const int qpos = m0 + t * kTcBM + row;           // query index inside this sequence's chunk
        const int limit = off + qpos;                    // last visible key
        const int first_limit = off + m0 + t * kTcBM;    // ... of the tile's first row (CTA-uniform per tile)
        const uint32_t lane_sel = static_cast<uint32_t>(quarter * 32) << 16;
        const uint32_t tmem_S = tmem_base + t * 128 + lane_sel;
        const uint32_t tmem_O = tmem_base + 256 + t * 128 + lane_sel;
        float m_run = -INFINITY, l_run = 0.f;
        for (int jj = 0; jj < 2 * nt; ++jj) {
            const int j = jj >> 1, hf = jj & 1;          // K/V tile, half of it
            const int kv0 = j * kTcBN + hf * kTcHalf;
            const int c_idx = t * 2 + hf;                // this hand-over chain
            mbar_wait(&s_full[c_idx], j & 1);
            tc_fence_after();
            uint32_t sv[4][16];                          // 64 scores: one TMEM round trip
#pragma unroll
            for (int c = 0; c < 4; ++c) tmem_ld16(tmem_S + hf * kTcHalf + c * 16, sv[c]);
            tmem_ld_wait();
            if (kv0 + kTcHalf - 1 > first_limit) {       // touches the causal boundary
#pragma unroll
                for (int c = 0; c < 4; ++c)
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                        if (kv0 + c * 16 + i > limit) sv[c][i] = 0xff800000u;  // -inf
            }
            float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};  // independent chains
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
                for (int i = 0; i < 16; ++i) mx4[i & 3] = fmaxf(mx4[i & 3], __uint_as_float(sv[c][i]));
            const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
            // Lazy reference update: the exponent reference only moves when some row of this warp saw its
            // maximum grow by more than 2^8 relative to it (P stays <= 256, exact in fp32 and fine in bf16; l
            // accumulates against the same reference, so the final O / l is unchanged). Rescaling O in TMEM is a
            // full round trip on the critical path and was taken on 97 % of tiles with an eager update.
            const float cand = mx * scale_log2;  // -inf when the row has seen nothing here
            float corr = 1.f;
            const bool moved = __any_sync(0xffffffffu, cand > m_run + 8.f);
            if (moved) {
                const float mnew = fmaxf(m_run, cand);
                corr = ex2_approx(m_run - mnew);   // first time: 2^-inf = 0
                m_run = mnew;
            }
            const float mref = m_run == -INFINITY ? 0.f : m_run;
            const f32x2 sc2 = pack2(scale_log2, scale_log2), nm2 = pack2(-mref, -mref);
            f32x2 ls2[2] = {pack2(0.f, 0.f), pack2(0.f, 0.f)};       // four independent sums
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
                for (int i = 0; i < 16; i += 2) {
                    float x0, x1;
                    unpack2(fma2(pack2(__uint_as_float(sv[c][i]), __uint_as_float(sv[c][i + 1])), sc2, nm2), x0, x1);
                    const float p0 = ex2_approx(x0), p1 = ex2_approx(x1);
                    ls2[(i >> 1) & 1] = add2(ls2[(i >> 1) & 1], pack2(p0, p1));
                    sv[c][i >> 1] = pack_bf16(p0, p1);
                }
            float la, lb, lc, ld;
            unpack2(ls2[0], la, lb);
            unpack2(ls2[1], lc, ld);
            l_run = l_run * corr + (la + lb) + (lc + ld);
            // P as bf16 pairs into the first 32 columns of the half just read (the A operand of P.V comes from
            // TMEM: nothing of P crosses shared memory). No wait is needed here: QK of this chain was issued after
            // its previous P.V, so that one has retired.
            {
                uint32_t pw[16];
#pragma unroll
                for (int qd = 0; qd < 2; ++qd) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        pw[i] = sv[2 * qd][i];
                        pw[8 + i] = sv[2 * qd + 1][i];
                    }
                    tmem_st16(tmem_S + hf * kTcHalf + qd * 16, pw);
                }
            }
            // rescale this row of O only if some row of the warp moved its maximum; the P.V of the OTHER half
            // (issued one hand-over ago) may still be accumulating into O: wait for it first
            if (jj > 0 && moved) {
                mbar_wait(&o_done[t * 2 + (hf ^ 1)], (hf ? j : j - 1) & 1);
                tc_fence_after();
#pragma unroll 1
                for (int c = 0; c < kTcD; c += 16) {
                    uint32_t v[16];
                    tmem_ld16(tmem_O + c, v);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) * corr);
                    tmem_st16(tmem_O + c, v);
                }
            }
            tmem_st_wait();
            // keys past the context in the last tile: their V rows are stale page contents or never-loaded
            // shared memory; P is 0 there but 0 * NaN would poison O, so clear those rows (both dim-atoms)
            if (kv0 + kTcHalf > ctx) {
                mbar_wait(&v_full[j % kVS], (j / kVS) & 1);
                const int valid = ctx - j * kTcBN;       // rows of the 128-key tile that hold real keys
                if ((row >> 6) == hf && row >= valid) {
                    uint8_t* vrow = sV + (j % kVS) * kKVTile + row * 128;
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
#pragma unroll
                        for (int a = 0; a < kAtoms; ++a)
                            *reinterpret_cast<uint4*>(vrow + a * kKVAtom + i * 16) = make_uint4(0, 0, 0, 0);
                    }
                }
                fence_proxy_async();   // generic-proxy smem writes (cleared V rows) -> visible to the tensor core
            }
            tc_fence_before();
            mbar_arrive(&p_full[c_idx]);
        }
        // ---- epilogue: O / l -> bf16, one row of D values per thread
        if (nt > 0) {
            mbar_wait(&o_done[t * 2 + 1], (nt - 1) & 1);   // the last P.V of this tile
            tc_fence_after();
            const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
            __nv_bfloat16* orow = out + (static_cast<long long>(q_begin + qpos) * Hq + h) * kTcD;
#pragma unroll 1
            for (int c = 0; c < kTcD; c += 16) {
                uint32_t v[16];
                tmem_ld16(tmem_O + c, v);
                tmem_ld_wait();
                if (qpos < n_q) {
                    uint4 a, b;
                    a.x = pack_bf16(__uint_as_float(v[0]) * inv, __uint_as_float(v[1]) * inv);
                    a.y = pack_bf16(__uint_as_float(v[2]) * inv, __uint_as_float(v[3]) * inv);
                    a.z = pack_bf16(__uint_as_float(v[4]) * inv, __uint_as_float(v[5]) * inv);
                    a.w = pack_bf16(__uint_as_float(v[6]) * inv, __uint_as_float(v[7]) * inv);
                    b.x = pack_bf16(__uint_as_float(v[8]) * inv, __uint_as_float(v[9]) * inv);
                    b.y = pack_bf16(__uint_as_float(v[10]) * inv, __uint_as_float(v[11]) * inv);
                    b.z = pack_bf16(__uint_as_float(v[12]) * inv, __uint_as_float(v[13]) * inv);
                    b.w = pack_bf16(__uint_as_float(v[14]) * inv, __uint_as_float(v[15]) * inv);
                    *reinterpret_cast<uint4*>(orow + c) = a;
                    *reinterpret_cast<uint4*>(orow + c + 8) = b;
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<512>(tmem_base);
}

// ---- host: cached TMA descriptors (the runner calls with the same buffers every step) ----------------------
typedef CUresult (*EncodeTiledFn2)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

struct MapKey {
    const void* base;
    long long rows, cols;
    int box_rows;
    CUtensorMap map;
};

static int cached_map(const void* base, long long rows, long long cols, int box_rows, CUtensorMap* out) {
    static std::mutex mu;
    static std::vector<MapKey> cache;
    static EncodeTiledFn2 fn = nullptr;
    std::lock_guard<std::mutex> lock(mu);
    for (const MapKey& k : cache)
        if (k.base == base && k.rows == rows && k.cols == cols && k.box_rows == box_rows) {
            *out = k.map;
            return ATX_OK;
        }
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess) {
            set_error("cuTensorMapEncodeTiled not available from the driver");
            return ATX_ERR_CUDA;
        }
        fn = reinterpret_cast<EncodeTiledFn2>(p);
    }
    MapKey k = {base, rows, cols, box_rows, {}};
    cuuint64_t dims[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
    cuuint64_t strides[1] = {static_cast<cuuint64_t>(cols) * 2};
    cuuint32_t box[2] = {64, static_cast<cuuint32_t>(box_rows)};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(&k.map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (CUresult %d) rows=%lld cols=%lld", static_cast<int>(r), rows, cols);
        return ATX_ERR_CUDA;
    }
    if (cache.size() >= 512) cache.erase(cache.begin());
    cache.push_back(k);
    *out = k.map;
    return ATX_OK;
}

template <int D>
static int launch_prefill_tc_d(const void* q, const void* kc, const void* vc, const int* bt, const int* cu_q,
                               const int* ctx, void* out, int S, int max_q_len, long long total_q, long long num_blocks,
                               int Hq, int Hkv, int max_blocks, float scale, int flags, cudaStream_t stream) {
    CUtensorMap tmQ, tmK, tmV;
    int rc = cached_map(q, total_q, static_cast<long long>(Hq) * D, kTcBM, &tmQ);
    if (rc == ATX_OK) rc = cached_map(kc, num_blocks * Hkv * kTcPage, D, kTcPage, &tmK);
    if (rc == ATX_OK) rc = cached_map(vc, num_blocks * Hkv * kTcPage, D, kTcPage, &tmV);
    if (rc != ATX_OK) return rc;
    static bool attr_done[64] = {false};
    int dev = 0;
    ATX_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !attr_done[dev]) {
        ATX_CUDA(cudaFuncSetAttribute(attn_prefill_tc_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      TcCfg<D>::kSmem));
        attr_done[dev] = true;
    }
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = (flags & ATX_FLAG_PDL) ? 1 : 0;
    const int n_mtiles = (max_q_len + 2 * kTcBM - 1) / (2 * kTcBM);
    const long long n_ctas = static_cast<long long>(n_mtiles) * Hq * S;
    if (n_ctas > 0x7fffffffll) {
        set_error("atx_paged_attn_prefill: %lld CTAs exceed the grid limit", n_ctas);
        return ATX_ERR_LIMIT;
    }
    cfg.gridDim = dim3(static_cast<unsigned>(n_ctas));
    cfg.blockDim = dim3(kTcThreads);
    cfg.dynamicSmemBytes = TcCfg<D>::kSmem;
    cfg.stream = stream;
    ATX_CUDA(cudaLaunchKernelEx(&cfg, attn_prefill_tc_kernel<D>, tmQ, tmK, tmV, bt, cu_q, ctx,
                                reinterpret_cast<__nv_bfloat16*>(out), Hq, Hkv, max_blocks, n_mtiles,
                                scale * 1.4426950408889634f));
    return ATX_OK;
}

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
    // One kernel, on tcgen05: TMA boxes are one 16-token page of one KV head, the swizzled tiles need 16-byte rows
    if ((D != 64 && D != 128) || block_size != kTcPage || (reinterpret_cast<uintptr_t>(q) & 15) != 0 ||
        (reinterpret_cast<uintptr_t>(k_cache) & 15) != 0 || (reinterpret_cast<uintptr_t>(v_cache) & 15) != 0) {
        set_error("atx_paged_attn_prefill: unsupported head_dim=%d (64 or 128) / block_size=%d (16) / unaligned buffers",
                  D, block_size);
        return ATX_ERR_UNSUPPORTED;
    }
    return D == 128 ? launch_prefill_tc_d<128>(q, k_cache, v_cache, block_table, cu_q, ctx_lens, out, S, max_q_len, total_q,
                                               num_blocks, Hq, Hkv, max_blocks, scale, flags, stream)
                    : launch_prefill_tc_d<64>(q, k_cache, v_cache, block_table, cu_q, ctx_lens, out, S, max_q_len, total_q,
                                              num_blocks, Hq, Hkv, max_blocks, scale, flags, stream);
}

}  // extern "C"
