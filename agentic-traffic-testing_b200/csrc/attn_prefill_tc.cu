// Causal prefill attention over the paged KV cache on tcgen05 tensor cores (head_dim 128, 16-token pages).
//
// Same arithmetic as attn_prefill.cu (transformers modeling_llama.py:199-222, fp32 softmax :217); both GEMMs of
// attention run on tcgen05 with TMEM accumulators:
//
//   S  = Q K^T   M=128 query rows, N=64 keys,  K=128 dims     A=Q (smem, K-major)  B=K  (smem, K-major)
//   O += P V     M=128 query rows, N=128 dims, K=64 keys      A=P (TMEM)           B=V  (smem, MN-major)
//
// One CTA = TWO 128-row query tiles (A: rows m0.., B: rows m0+128..) of one head, worked in ping-pong: while the
// four softmax warps of one tile are busy (TMEM -> registers, mask, max, ex2, P -> smem) the tensor core runs
// the other tile's QK^T / PV, so neither side idles on the other. (The first version had one tile per CTA with
// two threads per row; ncu showed the tensor pipe 31 % active and the softmax warps, with two block barriers
// and a shared-memory exchange per tile, as the critical path: profiles/r01n.) Roles:
//   warp 0      TMA producer: Q_A, Q_B once; K and V in 64-key tiles (four pages; a page of one KV head is
//               16 rows x 256 B, two 128B-swizzled boxes), 3-deep rings, K one tile ahead of V
//   warp 1      the single MMA-issuing thread: QK_A(j+1), QK_B(j+1), then PV_A(j), PV_B(j) as their P arrives
//   warps 2-5   softmax of tile A, warps 6-9 of tile B: ONE thread per query row (no cross-thread exchange):
//               64 scores from TMEM in one round trip, causal mask on diagonal tiles, running max with a lazy
//               reference (O in TMEM is rescaled only when some row of the warp grew by more than 2^8), ex2,
//               P as bf16 pairs written back into the first 32 columns of the S buffer just read (tcgen05.st):
//               P.V takes its A operand from TMEM, P never crosses shared memory; finally O / l -> bf16
// TMEM (512 columns): S_A x2, S_B x2 (64 columns each, double-buffered so QK^T(j+1) overlaps softmax(j)),
// O_A, O_B (128 each). K and V tiles are shared by both query tiles: half the K/V shared-memory traffic per row.
// Ablations on B200 (S=4 x 8192, 2.2 ms): dropping the exponentials, the maxima, the P hand-over or half of the
// TMEM reads changes nothing, dropping every K/V load after the first tiles gains 7 %, two threads per row (four
// softmax warps per scheduler) LOSES 8 %: what is left is the latency of the hand-overs between the tensor pipe and
// the softmax warps (commit -> mbarrier -> tcgen05.ld behind queued MMAs -> tcgen05.st -> arrive), not a throughput.
#include "common.cuh"

#include <mutex>
#include <vector>

#include "../../include/atx_b200.h"

namespace atx {

constexpr int kTcBM = 128;        // query rows per tile (two tiles per CTA)
constexpr int kTcBN = 64;         // keys per tile
constexpr int kTcD = 128;
constexpr int kTcThreads = 320;   // TMA warp, MMA warp, 4 + 4 softmax warps
constexpr int kTcPage = 16;
constexpr int kQAtom = 128 * 128;               // [128 rows x 64 bf16], 128B-swizzled: one K-atom of Q (or of P)
constexpr int kQTile = 2 * kQAtom;              // [128 x 128] bf16
constexpr int kKVAtom = kTcBN * 128;            // [64 rows x 64 bf16]
constexpr int kKVTile = 2 * kKVAtom;            // [64 keys x 128 dims] bf16 as two dim-atoms
constexpr int kKS = 3, kVS = 3;                 // ring depths
constexpr int kTcSmem = 2 * kQTile + (kKS + kVS) * kKVTile + 1024 + 512;   // + align + barriers

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

__global__ void __launch_bounds__(kTcThreads, 1)
attn_prefill_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                       const __grid_constant__ CUtensorMap tmV, const int* __restrict__ block_table,
                       const int* __restrict__ cu_q, const int* __restrict__ ctx_lens,
                       __nv_bfloat16* __restrict__ out, int Hq, int Hkv, int max_blocks, float scale_log2) {
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
    uint64_t* s_full = bars + 17;       // [tile][buffer]: 4
    uint64_t* p_full = bars + 21;       // [tile]: 2
    uint64_t* o_done = bars + 23;       // [tile]: 2
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 25);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int s = blockIdx.z, h = blockIdx.y;
    const int q_begin = cu_q[s];
    const int n_q = cu_q[s + 1] - q_begin;
    const int m0 = blockIdx.x * 2 * kTcBM;
    if (m0 >= n_q) return;  // whole CTA: before any barrier / TMEM use
    const int ctx = ctx_lens[s];
    const int off = ctx - n_q;  // query j sees keys [0, off + j]
    const int g = h / (Hq / Hkv);
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
        }
        for (int i = 0; i < 2; ++i) {
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
    // columns: S_A[0] 0, S_A[1] 64, S_B[0] 128, S_B[1] 192, O_A 256, O_B 384
    pdl_launch_dependents();

    if (warp == 0) {
        // ===== TMA producer: the whole warp works. Lane p < 4 owns page p of a tile: the block-table lookups of
        // a tile are one coalesced load, two tiles ahead of their use, and lane p issues its page's boxes.
        pdl_wait();  // q and this chunk's K/V rows come from the upstream kernels
        if (lane == 0) {
            mbar_expect_tx(q_full, has_b ? 2 * kQTile : kQTile);
            for (int t = 0; t < (has_b ? 2 : 1); ++t) {
                tma_load_2d(sQ + t * kQTile, &tmQ, q_full, h * kTcD, q_begin + m0 + t * kTcBM, kEvictNormal);
                tma_load_2d(sQ + t * kQTile + kQAtom, &tmQ, q_full, h * kTcD + 64, q_begin + m0 + t * kTcBM,
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
                mbar_expect_tx(&full[st], tile_pages(j) * kTcPage * 256);   // allocated pages only
            }
            __syncwarp();
            if (row >= 0) {
                uint8_t* dst = ring + st * kKVTile + lane * kTcPage * 128;
                tma_load_2d(dst, tm, &full[st], 0, row, kEvictLast);
                tma_load_2d(dst + kKVAtom, tm, &full[st], 64, row, kEvictLast);
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
            // ===== MMA issuer. tcgen05.mma execute in issue order, which is what makes the S buffers safe:
            // QK_X(j+1) overwrites S_X[(j+1)&1], last read by softmax_X(j-1), whose P the issuer has waited for.
            constexpr uint32_t idesc_qk = umma_idesc_bf16(kTcBM, kTcBN);                 // A, B K-major
            constexpr uint32_t idesc_pv = umma_idesc_bf16(kTcBM, kTcD) | (1u << 16);      // B (= V) MN-major
            auto issue_qk = [&](int j) {
                const int st = j % kKS;
                mbar_wait(&k_full[st], (j / kKS) & 1);
                tc_fence_after();
                const uint32_t ka = smem_u32(sK + st * kKVTile);
#pragma unroll
                for (int t = 0; t < 2; ++t) {
                    if (t == 0 ? j >= nt_a : j >= nt_b) continue;
                    const uint32_t qa = smem_u32(sQ + t * kQTile);
                    const uint32_t d_tmem = tmem_base + t * 128 + (j & 1) * kTcBN;
#pragma unroll
                    for (int kk = 0; kk < kTcD / 16; ++kk)
                        umma_bf16(d_tmem, umma_desc_sw128(qa + (kk >> 2) * kQAtom + (kk & 3) * 32),
                                  umma_desc_sw128(ka + (kk >> 2) * kKVAtom + (kk & 3) * 32), idesc_qk, kk > 0 ? 1u : 0u);
                    umma_commit(&s_full[t * 2 + (j & 1)]);
                }
                umma_commit(&k_empty[st]);
            };
            mbar_wait(q_full, 0);
            issue_qk(0);
            for (int j = 0; j < n_tiles; ++j) {
                if (j + 1 < n_tiles) issue_qk(j + 1);  // overlaps the softmax of tile j
                const int st = j % kVS;
                mbar_wait(&v_full[st], (j / kVS) & 1);
                const uint32_t va = smem_u32(sV + st * kKVTile);
#pragma unroll
                for (int t = 0; t < 2; ++t) {
                    if (t == 0 ? j >= nt_a : j >= nt_b) continue;
                    mbar_wait(&p_full[t], j & 1);          // P_t(j) in smem, O_t rescaled, tail rows of V cleared
                    tc_fence_after();
                    const uint32_t pa = tmem_base + t * 128 + (j & 1) * kTcBN;   // P_t(j) lives where S_t(j) was
#pragma unroll
                    for (int kk = 0; kk < kTcBN / 16; ++kk)
                        umma_bf16_ts(tmem_base + 256 + t * 128, pa + kk * 8,
                                     umma_desc_sw128_mn(va + kk * 16 * 128, kKVAtom), idesc_pv,
                                     (j > 0 || kk > 0) ? 1u : 0u);
                    umma_commit(&o_done[t]);
                }
                umma_commit(&v_empty[st]);
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
        for (int j = 0; j < nt; ++j) {
            const int kv0 = j * kTcBN;
            mbar_wait(&s_full[t * 2 + (j & 1)], (j >> 1) & 1);
            tc_fence_after();
            uint32_t sv[4][16];                          // 64 scores: one TMEM round trip
#pragma unroll
            for (int c = 0; c < 4; ++c) tmem_ld16(tmem_S + (j & 1) * kTcBN + c * 16, sv[c]);
            tmem_ld_wait();
            if (kv0 + kTcBN - 1 > first_limit) {         // tile touches the causal boundary
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
            const float cand = mx * scale_log2;  // -inf when the row has seen nothing in this tile
            float corr = 1.f;
            const bool moved = __any_sync(0xffffffffu, cand > m_run + 8.f);
            if (moved) {
                const float mnew = fmaxf(m_run, cand);
                corr = ex2_approx(m_run - mnew);   // first tile: 2^-inf = 0
                m_run = mnew;
            }
            const float mref = m_run == -INFINITY ? 0.f : m_run;
            float ls4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
                for (int i = 0; i < 16; i += 2) {
                    const float p0 = ex2_approx(fmaf(__uint_as_float(sv[c][i]), scale_log2, -mref));
                    const float p1 = ex2_approx(fmaf(__uint_as_float(sv[c][i + 1]), scale_log2, -mref));
                    ls4[(i >> 1) & 3] += p0 + p1;
                    sv[c][i >> 1] = pack_bf16(p0, p1);
                }
            l_run = l_run * corr + (ls4[0] + ls4[1]) + (ls4[2] + ls4[3]);
            // Stay at most one tile ahead of the tensor core: p_full is a one-bit-phase barrier, and a second
            // arrival round before the issuer has seen the first would leave it waiting for ever. (P.V(j-1) has
            // retired: also what the O rescale below needs.)
            if (j > 0) mbar_wait(&o_done[t], (j - 1) & 1);
            tc_fence_after();
            // P as bf16 pairs into the first 32 columns of the S buffer just read (the A operand of P.V comes
            // from TMEM: nothing of P crosses shared memory)
            {
                uint32_t pw[16];
#pragma unroll
                for (int hlf = 0; hlf < 2; ++hlf) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        pw[i] = sv[2 * hlf][i];
                        pw[8 + i] = sv[2 * hlf + 1][i];
                    }
                    tmem_st16(tmem_S + (j & 1) * kTcBN + hlf * 16, pw);
                }
            }
            // rescale this row of O only if some row of the warp moved its maximum
            if (j > 0 && moved) {
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
            if (kv0 + kTcBN > ctx) {
                mbar_wait(&v_full[j % kVS], (j / kVS) & 1);
                const int valid = ctx - kv0;
                if (row < kTcBN && row >= valid) {
                    uint8_t* vrow = sV + (j % kVS) * kKVTile + row * 128;
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        *reinterpret_cast<uint4*>(vrow + i * 16) = make_uint4(0, 0, 0, 0);
                        *reinterpret_cast<uint4*>(vrow + kKVAtom + i * 16) = make_uint4(0, 0, 0, 0);
                    }
                }
                fence_proxy_async();   // generic-proxy smem writes (cleared V rows) -> visible to the tensor core
            }
            tc_fence_before();
            mbar_arrive(&p_full[t]);
        }
        // ---- epilogue: O / l -> bf16, one 256-byte row per thread
        if (nt > 0) {
            mbar_wait(&o_done[t], (nt - 1) & 1);
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

int launch_prefill_tc(const void* q, const void* kc, const void* vc, const int* bt, const int* cu_q, const int* ctx,
                      void* out, int S, int max_q_len, long long total_q, long long num_blocks, int Hq, int Hkv,
                      int max_blocks, float scale, int flags, cudaStream_t stream) {
    CUtensorMap tmQ, tmK, tmV;
    int rc = cached_map(q, total_q, static_cast<long long>(Hq) * kTcD, kTcBM, &tmQ);
    if (rc == ATX_OK) rc = cached_map(kc, num_blocks * Hkv * kTcPage, kTcD, kTcPage, &tmK);
    if (rc == ATX_OK) rc = cached_map(vc, num_blocks * Hkv * kTcPage, kTcD, kTcPage, &tmV);
    if (rc != ATX_OK) return rc;
    static bool attr_done[64] = {false};
    int dev = 0;
    ATX_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !attr_done[dev]) {
        ATX_CUDA(cudaFuncSetAttribute(attn_prefill_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmem));
        attr_done[dev] = true;
    }
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = (flags & ATX_FLAG_PDL) ? 1 : 0;
    cfg.gridDim = dim3(static_cast<unsigned>((max_q_len + 2 * kTcBM - 1) / (2 * kTcBM)), static_cast<unsigned>(Hq),
                       static_cast<unsigned>(S));
    cfg.blockDim = dim3(kTcThreads);
    cfg.dynamicSmemBytes = kTcSmem;
    cfg.stream = stream;
    ATX_CUDA(cudaLaunchKernelEx(&cfg, attn_prefill_tc_kernel, tmQ, tmK, tmV, bt, cu_q, ctx,
                                reinterpret_cast<__nv_bfloat16*>(out), Hq, Hkv, max_blocks,
                                scale * 1.4426950408889634f));
    return ATX_OK;
}

}  // namespace atx
