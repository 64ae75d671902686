// bf16 projection / MLP / LM-head GEMM on tcgen05 tensor cores with TMEM accumulators.
//
//   out[t, n] (+)= sum_k X[t, k] * W[n, k]          X: [T, K] bf16 activations, W: [N, K] bf16
//
// This is torch.nn.functional.linear(x, W) as used by the reference's HF path
// (transformers/models/llama/modeling_llama.py:182-184 MLP, :259-289 q/k/v/o, :499 lm_head) and by
// vLLM's LlamaForCausalLM behind llm/serve_llm.py:527.
//
// B200 mapping ("swap-AB"): the WEIGHT matrix is the UMMA A operand (M = 128 output features per
// CTA, one TMEM lane each) and the TOKENS are the UMMA B operand (N = TN tokens, TN in 16..256).
// Decode batches (T = 1..32) therefore waste no tensor rows, and the kernel is what decode needs it
// to be: a TMA-fed weight streamer. Both operands are K-major, staged by TMA into 128B-swizzled
// shared memory through a kStages-deep mbarrier ring; one elected thread issues tcgen05.mma; the
// fp32 accumulator tile [128 x TN] lives in TMEM and is read back with tcgen05.ld for the epilogue.
// Split-K partial sums are combined with fp32 red.global.add straight into the destination, which
// for o_proj / down_proj IS the fp32 residual stream (residual add fused, no workspace).
#include "common.cuh"

#include <cstdlib>
#include <mutex>
#include <new>

#include "../../include/atx_b200.h"

namespace atx {

constexpr int kBM = 128;  // weight rows per CTA (UMMA M)
constexpr int kBK = 64;   // K elements per stage (= one 128-byte swizzle row of bf16)
constexpr int kUK = 16;   // UMMA K for 16-bit inputs
constexpr int kGemmThreads = 128;

template <int TN>
struct GemmCfg {
    static constexpr int kStageBytesW = kBM * kBK * 2;  // 16 KB
    static constexpr int kStageBytesX = TN * kBK * 2;
    static constexpr int kStageBytes = kStageBytesW + kStageBytesX;
    // decode tiles stay small enough for several CTAs per SM; prefill tiles take most of the SM
#ifndef ATX_STAGES_SMALL
#define ATX_STAGES_SMALL 6
#endif
    // TN=16: 6 x 18 KB and TN=32: 5 x 20 KB both leave room for two CTAs per SM
    static constexpr int kStages = TN <= 16 ? ATX_STAGES_SMALL : (TN <= 32 ? 5 : (TN <= 64 ? 6 : (TN <= 128 ? 6 : 4)));
    static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;
    static constexpr uint32_t kTmemCols = TN < 32 ? 32 : TN;
};

__device__ __forceinline__ void cluster_barrier() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t dsmem_addr(const void* local_ptr, uint32_t cta_rank) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(local_ptr)), "r"(cta_rank));
    return remote;
}
__device__ __forceinline__ float dsmem_ld(uint32_t remote_addr) {
    float v;
    asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(remote_addr) : "memory");
    return v;
}

template <int TN>
__global__ void __launch_bounds__(kGemmThreads)
gemm_wt_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmX,
               void* __restrict__ out, long long ldo, int T, int N, int kb_total, int kb_per_split,
               int mode_in, int w_evict_first, int stages) {
    const int mode = mode_in & 0xff;
    const bool ordered = (mode_in & 0x100) != 0;   // ATX_FLAG_ORDERED: split-K summed in a fixed order (below)
    // `stages` (<= Cfg::kStages) is the ring depth actually used: CTAs with few k-blocks get a smaller
    // shared-memory footprint, so more of them are resident per SM
    using Cfg = GemmCfg<TN>;
    extern __shared__ uint8_t smem_raw[];
    // SWIZZLE_128B tiles must sit on 1024-byte boundaries
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + stages * Cfg::kStageBytes);
    uint64_t* empty_bar = full_bar + Cfg::kStages;
    uint64_t* accum_bar = empty_bar + Cfg::kStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int n0 = blockIdx.x * kBM;
    const int t0 = blockIdx.y * TN;
    const int kb_begin = blockIdx.z * kb_per_split;
    const int kb_end = min(kb_total, kb_begin + kb_per_split);
    const int nkb = kb_end - kb_begin;

    if (warp == 0 && elect_one()) {
        tma_prefetch_desc(&tmW);
        tma_prefetch_desc(&tmX);
#pragma unroll
        for (int s = 0; s < Cfg::kStages; ++s) {
            if (s >= stages) break;
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(accum_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_acc = *tmem_slot;
    // Let the next kernel in the stream start its own prologue (and, if it is a GEMM, its weight
    // prefetch) now; it still waits for this grid to finish before touching activations.
    pdl_launch_dependents();

    if (nkb > 0) {
        if (warp == 0) {
            if (elect_one()) {
                // ===== TMA producer. Weights never depend on the previous kernel, so with
                // programmatic dependent launch the first weight stages are already in flight
                // while the producer of X is still finishing; only X waits for it.
                const uint64_t polW = w_evict_first ? kEvictFirst : kEvictNormal;
                const int pre = min(nkb, stages);
                for (int i = 0; i < pre; ++i) {
                    mbar_expect_tx(&full_bar[i], Cfg::kStageBytes);
                    tma_load_2d(smem + i * Cfg::kStageBytes, &tmW, &full_bar[i],
                                (kb_begin + i) * kBK, n0, polW);
                }
                pdl_wait();
                for (int i = 0; i < pre; ++i) {
                    tma_load_2d(smem + i * Cfg::kStageBytes + Cfg::kStageBytesW, &tmX, &full_bar[i],
                                (kb_begin + i) * kBK, t0, kEvictLast);
                }
                int stage = 0;
                uint32_t phase = 0;
                for (int i = pre; i < nkb; ++i) {
                    mbar_wait(&empty_bar[stage], phase);
                    uint8_t* st = smem + stage * Cfg::kStageBytes;
                    mbar_expect_tx(&full_bar[stage], Cfg::kStageBytes);
                    tma_load_2d(st, &tmW, &full_bar[stage], (kb_begin + i) * kBK, n0, polW);
                    tma_load_2d(st + Cfg::kStageBytesW, &tmX, &full_bar[stage], (kb_begin + i) * kBK,
                                t0, kEvictLast);
                    if (++stage == stages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        } else if (warp == 1) {
            if (elect_one()) {
                // ===== MMA issuer (single thread)
                constexpr uint32_t idesc = umma_idesc_bf16(kBM, TN);
                int stage = 0;
                uint32_t phase = 0;
                for (int i = 0; i < nkb; ++i) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(smem + stage * Cfg::kStageBytes);
                    const uint32_t b_addr = a_addr + Cfg::kStageBytesW;
                    const uint64_t a_desc = umma_desc_sw128(a_addr);
                    const uint64_t b_desc = umma_desc_sw128(b_addr);
#pragma unroll
                    for (int k = 0; k < kBK / kUK; ++k) {
                        // advance 16 elements (32 B) along K inside the swizzle row: +2 in the
                        // 16-byte-granular start-address field
                        umma_bf16(tmem_acc, a_desc + 2 * k, b_desc + 2 * k, idesc,
                                  (i > 0 || k > 0) ? 1u : 0u);
                    }
                    umma_commit(&empty_bar[stage]);  // frees the smem slot when the MMAs retire
                    if (++stage == stages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                umma_commit(accum_bar);  // accumulator complete
            }
        }
        __syncwarp();

        // ===== epilogue: all four warps; warp w owns TMEM lanes [32w, 32w+32) = weight rows
        mbar_wait(accum_bar, 0);
        tc_fence_after();
        pdl_wait();  // `out` may be the residual stream written by the upstream grid
        const bool nonlinear = mode == ATX_GEMM_SWIGLU_BF16 || ordered;
        if (nonlinear && gridDim.z > 1) {
            // Split-K with a nonlinear epilogue (or with ATX_FLAG_ORDERED, where the order of the additions must
            // not depend on timing): the gridDim.z CTAs of this tile form ONE thread-block
            // cluster (launch attribute). CTAs 1.. park their partial accumulator in their own (idle) stage
            // buffers, the leader pulls them through distributed shared memory in a fixed order, folds
            // them into its TMEM accumulator and runs the ordinary SwiGLU epilogue: no workspace, no
            // atomics, bit-reproducible sums. The second cluster barrier keeps the peers' shared memory
            // alive until the leader has read it.
            float* part = reinterpret_cast<float*>(smem);  // [TN][128]
            const int row = warp * 32 + lane;
            const uint32_t tl = tmem_acc + (static_cast<uint32_t>(warp * 32) << 16);
            if (blockIdx.z != 0) {
#pragma unroll 1
                for (int c = 0; c < TN; c += 16) {
                    if (t0 + c >= T) break;
                    uint32_t v[16];
                    tmem_ld16(tl + c, v);
                    tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 16; ++j) part[(c + j) * kBM + row] = __uint_as_float(v[j]);
                }
            }
            cluster_barrier();
            if (blockIdx.z == 0) {
#pragma unroll 1
                for (int c = 0; c < TN; c += 16) {
                    if (t0 + c >= T) break;
                    uint32_t v[16];
                    tmem_ld16(tl + c, v);
                    tmem_ld_wait();
                    for (uint32_t r = 1; r < gridDim.z; ++r) {
                        const uint32_t remote = dsmem_addr(part + c * kBM + row, r);
#pragma unroll
                        for (int j = 0; j < 16; ++j)
                            v[j] = __float_as_uint(__uint_as_float(v[j]) + dsmem_ld(remote + j * kBM * 4));
                    }
                    tmem_st16(tl + c, v);
                }
                tmem_st_wait();
            }
            cluster_barrier();
        }
        if (nonlinear && blockIdx.z != 0) {
            // cluster peer: its sum now lives in the leader's accumulator
        } else if (mode == ATX_GEMM_SWIGLU_BF16) {
            // Rows [0,64) of this tile are gate features, rows [64,128) the matching up features
            // (weights interleaved in 64-row groups at load time): act = silu(gate) * up, bf16.
            // The halves cross between warps 0-1 (gate lanes) and warps 2-3 (up lanes) through the
            // (now idle) stage buffers. For TN >= 32 the exchange is symmetric so all four warps do
            // the activation math: warps 0-1 finish token columns [0, TN/2), warps 2-3 the rest.
            float* sx = reinterpret_cast<float*>(smem);  // [TN][64]: up for cols < HALF, gate for cols >= HALF
            constexpr int HALF = TN >= 32 ? TN / 2 : TN;
            const bool gate_warp = warp < 2;
            const int i = (warp & 1) * 32 + lane;            // feature within the 64-group
            const long long col = static_cast<long long>(blockIdx.x) * 64 + i;
            const bool ok = col < N / 2;
            __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out);
            const uint32_t tlane = tmem_acc + (static_cast<uint32_t>(warp * 32) << 16);
            // phase 1: park the half the OTHER pair of warps needs
            {
                const int c_lo = gate_warp ? HALF : 0, c_hi = gate_warp ? TN : HALF;
#pragma unroll 1
                for (int c = c_lo; c < c_hi; c += 16) {
                    if (t0 + c >= T) break;
                    uint32_t v[16];
                    tmem_ld16(tlane + c, v);
                    tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 16; ++j) sx[(c + j) * 64 + i] = __uint_as_float(v[j]);
                }
            }
            __syncthreads();
            // phase 2: own half from TMEM, partner half from shared memory
            {
                const int c_lo = gate_warp ? 0 : HALF, c_hi = gate_warp ? HALF : TN;
#pragma unroll 1
                for (int c = c_lo; c < c_hi; c += 16) {
                    if (t0 + c >= T) break;
                    uint32_t v[16];
                    tmem_ld16(tlane + c, v);
                    tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const int t = t0 + c + j;
                        const float mine = __uint_as_float(v[j]), other = sx[(c + j) * 64 + i];
                        const float g = gate_warp ? mine : other, u = gate_warp ? other : mine;
                        if (ok && t < T)
                            o[static_cast<long long>(t) * ldo + col] = __float2bfloat16_rn(g / (1.f + expf(-g)) * u);
                    }
                }
            }
        } else {
        const int n = n0 + warp * 32 + lane;
        const bool n_ok = n < N;
#pragma unroll 1
        for (int c = 0; c < TN; c += 16) {
            if (t0 + c >= T) break;  // warp-uniform
            uint32_t v[16];
            tmem_ld16(tmem_acc + (static_cast<uint32_t>(warp * 32) << 16) + c, v);
            tmem_ld_wait();
            if (n_ok) {
                if (mode == ATX_GEMM_STORE_F32) {
                    float* o = reinterpret_cast<float*>(out);
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const int t = t0 + c + j;
                        if (t < T) o[static_cast<long long>(t) * ldo + n] = __uint_as_float(v[j]);
                    }
                } else if (mode == ATX_GEMM_ADD_F32) {
                    // parked in shared memory (token-major) and added below, four consecutive outputs at a time
                    float* part = reinterpret_cast<float*>(smem);  // [TN][128]: the stage buffers are idle now
#pragma unroll
                    for (int j = 0; j < 16; ++j) part[(c + j) * kBM + warp * 32 + lane] = __uint_as_float(v[j]);
                } else {
                    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out);
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const int t = t0 + c + j;
                        if (t < T)
                            o[static_cast<long long>(t) * ldo + n] = __float2bfloat16_rn(__uint_as_float(v[j]));
                    }
                }
            }
        }
        if (mode == ATX_GEMM_ADD_F32) {
            // The split-K / residual add as 16-byte reductions: a thread's accumulator lane is ONE output column, so
            // added straight from registers every value is its own red.add (with T = 32 and 26 K-splits the qkv GEMM
            // of a 70B tensor-parallel shard issues a million of them onto 40 K addresses: the L2's atomic units,
            // not HBM, set its time). Through shared memory a thread owns four consecutive columns of a token.
            __syncthreads();
            const float* part = reinterpret_cast<const float*>(smem);
            float* o = reinterpret_cast<float*>(out);
            const int rows = min(TN, T - t0);
            const bool vec = (N & 3) == 0 && (ldo & 3) == 0 && (reinterpret_cast<uintptr_t>(o) & 15) == 0;
            for (int idx = threadIdx.x; idx < rows * (kBM / 4); idx += kGemmThreads) {
                const int tt = idx / (kBM / 4), n4 = (idx % (kBM / 4)) * 4;
                const int nn = n0 + n4;
                if (nn >= N) continue;
                const float4 a = *reinterpret_cast<const float4*>(part + tt * kBM + n4);
                float* dst = o + static_cast<long long>(t0 + tt) * ldo + nn;
                if (vec) {
                    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(a.x), "f"(a.y), "f"(a.z),
                                 "f"(a.w)
                                 : "memory");
                } else {
                    atomicAdd(dst, a.x);
                    if (nn + 1 < N) atomicAdd(dst + 1, a.y);
                    if (nn + 2 < N) atomicAdd(dst + 2, a.z);
                    if (nn + 3 < N) atomicAdd(dst + 3, a.w);
                }
            }
        }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<Cfg::kTmemCols>(tmem_acc);
}

// ---- persistent, warp-specialised variant for prefill-sized token tiles ----------------------------
// One CTA per SM loops over output tiles; a TMA producer warp, a single-thread MMA issuer and four
// epilogue warps run as three decoupled pipelines: the smem ring (full/empty mbarriers) between TMA and
// MMA, and TWO TMEM accumulators (tmem_full/tmem_empty) between MMA and epilogue, so the epilogue of
// tile i (TMEM -> registers -> global) overlaps the main loop of tile i+1 instead of idling the tensor
// pipe. Tiles are ordered token-tile-fastest so that CTAs running concurrently share weight tiles in L2.
constexpr int kPersistThreads = 192;
constexpr int kXchgBytes = 16 * 1024;  // SwiGLU half-tile exchange, 32 token columns at a time

template <int TN>
struct PersistCfg {
    static constexpr int kStageBytesW = kBM * kBK * 2;
    static constexpr int kStageBytesX = TN * kBK * 2;
    static constexpr int kStageBytes = kStageBytesW + kStageBytesX;
    static constexpr int kStages = TN <= 128 ? 6 : 4;
    static constexpr int kSmemBytes = kStages * kStageBytes + kXchgBytes + 1024 + 256;
    static constexpr uint32_t kTmemCols = 2 * TN;  // two accumulators
};

__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

template <int TN>
__global__ void __launch_bounds__(kPersistThreads, 1)
gemm_wt_persistent_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmX,
                          void* __restrict__ out, long long ldo, int T, int N, int kb_total, int tiles_t,
                          int total_tiles, int mode, int w_evict_first) {
    using Cfg = PersistCfg<TN>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    float* xchg = reinterpret_cast<float*>(smem + Cfg::kStages * Cfg::kStageBytes);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes + kXchgBytes);
    uint64_t* empty_bar = full_bar + Cfg::kStages;
    uint64_t* tmem_full = empty_bar + Cfg::kStages;   // [2]
    uint64_t* tmem_empty = tmem_full + 2;             // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && elect_one()) {
        tma_prefetch_desc(&tmW);
        tma_prefetch_desc(&tmX);
#pragma unroll
        for (int s = 0; s < Cfg::kStages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&tmem_full[0], 1);
        mbar_init(&tmem_full[1], 1);
        mbar_init(&tmem_empty[0], 4);  // one arrival per epilogue warp
        mbar_init(&tmem_empty[1], 4);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_launch_dependents();

    if (warp == 0) {
        if (elect_one()) {
            // ===== TMA producer
            const uint64_t polW = w_evict_first ? kEvictFirst : kEvictNormal;
            int stage = 0;
            uint32_t phase = 0;
            bool waited = false;
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
                const int n0 = (tile / tiles_t) * kBM, t0 = (tile % tiles_t) * TN;
                for (int kb = 0; kb < kb_total; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* st = smem + stage * Cfg::kStageBytes;
                    mbar_expect_tx(&full_bar[stage], Cfg::kStageBytes);
                    tma_load_2d(st, &tmW, &full_bar[stage], kb * kBK, n0, polW);
                    if (!waited) {  // weights never depend on the upstream kernel, activations do
                        pdl_wait();
                        waited = true;
                    }
                    tma_load_2d(st + Cfg::kStageBytesW, &tmX, &full_bar[stage], kb * kBK, t0, kEvictLast);
                    if (++stage == Cfg::kStages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {
            // ===== MMA issuer
            constexpr uint32_t idesc = umma_idesc_bf16(kBM, TN);
            int stage = 0, acc = 0;
            uint32_t phase = 0, acc_phase = 0;
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
                mbar_wait(&tmem_empty[acc], acc_phase ^ 1);  // epilogue has drained this accumulator
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * TN;
                for (int kb = 0; kb < kb_total; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(smem + stage * Cfg::kStageBytes);
                    const uint64_t a_desc = umma_desc_sw128(a_addr);
                    const uint64_t b_desc = umma_desc_sw128(a_addr + Cfg::kStageBytesW);
#pragma unroll
                    for (int k = 0; k < kBK / kUK; ++k)
                        umma_bf16(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (kb > 0 || k > 0) ? 1u : 0u);
                    umma_commit(&empty_bar[stage]);
                    if (++stage == Cfg::kStages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                umma_commit(&tmem_full[acc]);
                if (++acc == 2) {
                    acc = 0;
                    acc_phase ^= 1;
                }
            }
        }
    } else {
        // ===== epilogue warps 2..5: TMEM lane quarter = warp % 4
        const int quarter = warp & 3;
        const bool gate_warp = quarter < 2;            // SwiGLU: lanes 0-63 gate rows, 64-127 up rows
        const int feat = (quarter & 1) * 32 + lane;    // feature within the 64-group (SwiGLU)
        int acc = 0;
        uint32_t acc_phase = 0;
        bool waited = false;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
            const int n0 = (tile / tiles_t) * kBM, t0 = (tile % tiles_t) * TN;
            mbar_wait(&tmem_full[acc], acc_phase);
            tc_fence_after();
            if (!waited) {  // `out` may be the residual stream written by the upstream grid
                pdl_wait();
                waited = true;
            }
            const uint32_t tlane = tmem_base + acc * TN + (static_cast<uint32_t>(quarter * 32) << 16);
            if (mode == ATX_GEMM_SWIGLU_BF16) {
                // symmetric exchange in groups of 32 token columns: gate warps finish columns [0, TN/2),
                // up warps columns [TN/2, TN); each parks the half the other pair needs
                constexpr int HALF = TN / 2;
                const long long col = static_cast<long long>(tile / tiles_t) * 64 + feat;
                const bool ok = col < N / 2;
                __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out);
                float* park_mine = xchg + (gate_warp ? 0 : 32 * 64);    // [32][64] floats each
                float* park_other = xchg + (gate_warp ? 32 * 64 : 0);
#pragma unroll 1
                for (int g0 = 0; g0 < HALF; g0 += 32) {
                    const int c_park = (gate_warp ? HALF : 0) + g0;   // columns whose value the partner needs
                    const int c_own = (gate_warp ? 0 : HALF) + g0;    // columns this warp finishes
#pragma unroll
                    for (int c = 0; c < 32; c += 16) {
                        uint32_t v[16];
                        tmem_ld16(tlane + c_park + c, v);
                        tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 16; ++j) park_mine[(c + j) * 64 + feat] = __uint_as_float(v[j]);
                    }
                    epi_bar_sync();
#pragma unroll
                    for (int c = 0; c < 32; c += 16) {
                        uint32_t v[16];
                        tmem_ld16(tlane + c_own + c, v);
                        tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            const int t = t0 + c_own + c + j;
                            const float mine = __uint_as_float(v[j]), other = park_other[(c + j) * 64 + feat];
                            const float gg = gate_warp ? mine : other, uu = gate_warp ? other : mine;
                            if (ok && t < T)
                                o[static_cast<long long>(t) * ldo + col] =
                                    __float2bfloat16_rn(gg / (1.f + expf(-gg)) * uu);
                        }
                    }
                    epi_bar_sync();  // the exchange buffer is reused by the next group
                }
            } else {
                const int n = n0 + quarter * 32 + lane;
                const bool n_ok = n < N;
#pragma unroll 1
                for (int c = 0; c < TN; c += 16) {
                    if (t0 + c >= T) break;
                    uint32_t v[16];
                    tmem_ld16(tlane + c, v);
                    tmem_ld_wait();
                    if (n_ok) {
                        if (mode == ATX_GEMM_STORE_F32) {
                            float* o = reinterpret_cast<float*>(out);
#pragma unroll
                            for (int j = 0; j < 16; ++j) {
                                const int t = t0 + c + j;
                                if (t < T) o[static_cast<long long>(t) * ldo + n] = __uint_as_float(v[j]);
                            }
                        } else if (mode == ATX_GEMM_ADD_F32) {
                            float* o = reinterpret_cast<float*>(out);
#pragma unroll
                            for (int j = 0; j < 16; ++j) {
                                const int t = t0 + c + j;
                                if (t < T) atomicAdd(o + static_cast<long long>(t) * ldo + n, __uint_as_float(v[j]));
                            }
                        } else {
                            __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out);
#pragma unroll
                            for (int j = 0; j < 16; ++j) {
                                const int t = t0 + c + j;
                                if (t < T)
                                    o[static_cast<long long>(t) * ldo + n] = __float2bfloat16_rn(__uint_as_float(v[j]));
                            }
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty[acc]);  // accumulator free for the tile after next
            if (++acc == 2) {
                acc = 0;
                acc_phase ^= 1;
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<Cfg::kTmemCols>(tmem_base);
}

// ---- host side ---------------------------------------------------------------------------------------

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}

// rows x cols bf16 row-major (row stride ld elements), box = box_rows x 64 columns, 128B swizzle
static int make_map(CUtensorMap* m, const void* base, long long rows, long long cols, long long ld,
                    int box_rows) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) {
        set_error("cuTensorMapEncodeTiled not available from the driver");
        return ATX_ERR_CUDA;
    }
    cuuint64_t dims[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
    cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * 2};
    cuuint32_t box[2] = {static_cast<cuuint32_t>(kBK), static_cast<cuuint32_t>(box_rows)};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box,
                    estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (CUresult %d) rows=%lld cols=%lld ld=%lld box_rows=%d",
                  static_cast<int>(r), rows, cols, ld, box_rows);
        return ATX_ERR_CUDA;
    }
    return ATX_OK;
}

static const int kTileSizes[5] = {16, 32, 64, 128, 256};

}  // namespace atx

struct atx_gemm_plan {
    CUtensorMap tmW;
    CUtensorMap tmX[5];  // one per token-tile size
    long long N, K, T_cap;
    int w_evict_first;
};

using namespace atx;

template <int TN>
static int launch_gemm(const atx_gemm_plan* p, int ti, void* out, long long ldo, int T, int mode,
                       int splits, int pdl, cudaStream_t stream) {
    using Cfg = GemmCfg<TN>;
    static bool attr_done[64] = {false};  // per instantiation and device; benign race (idempotent call)
    int dev = 0;
    ATX_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !attr_done[dev]) {
        ATX_CUDA(cudaFuncSetAttribute(gemm_wt_kernel<TN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      Cfg::kSmemBytes));
        attr_done[dev] = true;
    }
    const int kb_total = static_cast<int>(p->K / kBK);
    const int kb_per = (kb_total + splits - 1) / splits;
    const int zs = (kb_total + kb_per - 1) / kb_per;
    int stages = kb_per < Cfg::kStages ? kb_per : Cfg::kStages;
    // the fused-SwiGLU epilogue parks half a tile in the stage buffers: keep them large enough
    // (a whole fp32 tile when the split-K cluster parks partial sums there)
    const bool ordered = (mode & 0x100) != 0;
    if ((mode & 0xff) == ATX_GEMM_SWIGLU_BF16 || (mode & 0xff) == ATX_GEMM_ADD_F32) {
        // epilogues that park the tile in the (idle) stage buffers: half a tile for SwiGLU, a whole fp32 tile for the
        // split-K cluster reduction and for the 16-byte adds
        const bool whole = zs > 1 || (mode & 0xff) == ATX_GEMM_ADD_F32;
        const int need = (TN * (whole ? kBM : 64) * 4 + Cfg::kStageBytes - 1) / Cfg::kStageBytes;
        if (stages < need) stages = need;
    }
    const int smem_bytes = stages * Cfg::kStageBytes + 1024 + 256;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(static_cast<unsigned>((p->N + kBM - 1) / kBM), static_cast<unsigned>((T + TN - 1) / TN),
                       static_cast<unsigned>(zs));
    cfg.blockDim = dim3(kGemmThreads);
    cfg.dynamicSmemBytes = smem_bytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    int na = 0;
    if (pdl) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    if (((mode & 0xff) == ATX_GEMM_SWIGLU_BF16 || ordered) && zs > 1) {  // the K-splits of one tile are one cluster (DSMEM reduction)
        attr[na].id = cudaLaunchAttributeClusterDimension;
        attr[na].val.clusterDim.x = 1;
        attr[na].val.clusterDim.y = 1;
        attr[na].val.clusterDim.z = static_cast<unsigned>(zs);
        ++na;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    ATX_CUDA(cudaLaunchKernelEx(&cfg, gemm_wt_kernel<TN>, p->tmW, p->tmX[ti], out, ldo, T,
                                static_cast<int>(p->N), kb_total, kb_per, mode, p->w_evict_first, stages));
    return ATX_OK;
}

template <int TN>
static int launch_gemm_persistent(const atx_gemm_plan* p, int ti, void* out, long long ldo, int T, int mode, int pdl,
                                  cudaStream_t stream) {
    using Cfg = PersistCfg<TN>;
    static bool attr_done[64] = {false};
    int dev = 0;
    ATX_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !attr_done[dev]) {
        ATX_CUDA(cudaFuncSetAttribute(gemm_wt_persistent_kernel<TN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      Cfg::kSmemBytes));
        attr_done[dev] = true;
    }
    const int kb_total = static_cast<int>(p->K / kBK);
    const int tiles_n = static_cast<int>((p->N + kBM - 1) / kBM);
    const int tiles_t = (T + TN - 1) / TN;
    const int total = tiles_n * tiles_t;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(static_cast<unsigned>(total < num_sms() ? total : num_sms()));
    cfg.blockDim = dim3(kPersistThreads);
    cfg.dynamicSmemBytes = Cfg::kSmemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    ATX_CUDA(cudaLaunchKernelEx(&cfg, gemm_wt_persistent_kernel<TN>, p->tmW, p->tmX[ti], out, ldo, T,
                                static_cast<int>(p->N), kb_total, tiles_t, total, mode, p->w_evict_first));
    return ATX_OK;
}

extern "C" {

int atx_gemm_plan_create(atx_gemm_plan** plan, const void* W, long long N, long long K, const void* X,
                         long long T_cap, long long ldx) {
    ATX_REQUIRE(plan && W && X, "atx_gemm_plan_create: null pointer");
    ATX_REQUIRE(N > 0 && K > 0 && T_cap > 0 && ldx >= K, "atx_gemm_plan_create: bad shape N=%lld K=%lld T_cap=%lld ldx=%lld", N, K, T_cap, ldx);
    ATX_REQUIRE(K % kBK == 0, "atx_gemm_plan_create: K=%lld must be a multiple of %d", K, kBK);
    ATX_REQUIRE((reinterpret_cast<uintptr_t>(W) & 15) == 0 && (reinterpret_cast<uintptr_t>(X) & 15) == 0 &&
                    (ldx * 2) % 16 == 0,
                "atx_gemm_plan_create: W, X and the X row pitch must be 16-byte aligned");
    atx_gemm_plan* p = new (std::nothrow) atx_gemm_plan();
    ATX_REQUIRE(p != nullptr, "atx_gemm_plan_create: out of host memory");
    p->N = N;
    p->K = K;
    p->T_cap = T_cap;
    // weights larger than L2 are streamed once per step: do not let them evict activations / KV
    p->w_evict_first = (N * K * 2 > (64ll << 20)) ? 1 : 0;
    int rc = make_map(&p->tmW, W, N, K, K, kBM);
    for (int i = 0; i < 5 && rc == ATX_OK; ++i) rc = make_map(&p->tmX[i], X, T_cap, K, ldx, kTileSizes[i]);
    if (rc != ATX_OK) {
        delete p;
        return rc;
    }
    *plan = p;
    return ATX_OK;
}

void atx_gemm_plan_destroy(atx_gemm_plan* plan) { delete plan; }

int atx_gemm_auto_splits(long long N, long long K, int T) {
    // Decode-sized batches are pure weight streaming. Two CTAs fit per SM (shared memory), so cut K
    // until the grid fills ONE wave of 2 x SMs CTAs (no tail wave), keeping >= 4 k-blocks per CTA.
    if (T > 64) return 1;
    const long long tiles = (N + kBM - 1) / kBM;
    const int kb = static_cast<int>(K / kBK);
    const long long slots = 2ll * num_sms();
    if (tiles >= slots) return 1;
    int s = static_cast<int>(slots / tiles);
    if (s > kb / 4) s = kb / 4;
    if (s < 1) s = 1;
    // make the split even: ceil(kb / s) blocks each, drop empty trailing splits
    const int per = (kb + s - 1) / s;
    return (kb + per - 1) / per;
}

int atx_gemm_run(const atx_gemm_plan* p, void* out, long long ldo, int T, int mode, int splits, int flags,
                 atx_stream_t stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    ATX_REQUIRE(p && out, "atx_gemm_run: null pointer");
    ATX_REQUIRE(T > 0 && T <= p->T_cap, "atx_gemm_run: T=%d outside (0, %lld]", T, p->T_cap);
    ATX_REQUIRE(mode >= ATX_GEMM_STORE_F32 && mode <= ATX_GEMM_SWIGLU_BF16, "atx_gemm_run: bad mode %d", mode);
    ATX_REQUIRE(ldo >= (mode == ATX_GEMM_SWIGLU_BF16 ? p->N / 2 : p->N), "atx_gemm_run: ldo=%lld too small for N=%lld", ldo, p->N);
    ATX_REQUIRE(mode != ATX_GEMM_SWIGLU_BF16 || p->N % 128 == 0, "atx_gemm_run: SWIGLU mode needs N %% 128 == 0");
    // ATX_FLAG_ORDERED: the K-splits of an ADD are reduced inside a cluster in a fixed order (as for SwiGLU) and ONE CTA
    // adds each output element, instead of every split racing its red.add: bit-reproducible, a little slower
    const bool ordered = (flags & ATX_FLAG_ORDERED) && mode == ATX_GEMM_ADD_F32 && T <= 64;
    if (ordered && splits != 1) {
        const int want = splits > 0 ? splits : atx_gemm_auto_splits(p->N, p->K, T);
        splits = 1;
        while (splits * 2 <= want && splits * 2 <= 8 && (p->K / kBK) % (splits * 2) == 0) splits *= 2;
    }
    if (splits <= 0) {
        splits = 1;
        if (mode == ATX_GEMM_ADD_F32) splits = atx_gemm_auto_splits(p->N, p->K, T);
        if (mode == ATX_GEMM_SWIGLU_BF16 && T <= 64) {  // cluster split-K: a power of two, at most 8 CTAs
            const int want = atx_gemm_auto_splits(p->N, p->K, T);
            while (splits * 2 <= want && splits * 2 <= 8 && (p->K / kBK) % (splits * 2) == 0) splits *= 2;
        }
    }
    ATX_REQUIRE(splits == 1 || mode == ATX_GEMM_ADD_F32 || mode == ATX_GEMM_SWIGLU_BF16,
                "atx_gemm_run: split-K (%d) needs ATX_GEMM_ADD_F32 into a zeroed or residual buffer, or "
                "ATX_GEMM_SWIGLU_BF16 (cluster reduction)", splits);
    ATX_REQUIRE(mode != ATX_GEMM_SWIGLU_BF16 || splits == 1 ||
                    (T <= 64 && splits <= 8 && (splits & (splits - 1)) == 0 && (p->K / kBK) % splits == 0),
                "atx_gemm_run: SWIGLU split-K needs T <= 64 and splits in {2,4,8} dividing K/64 (got T=%d splits=%d)",
                T, splits);
    const int pdl = (flags & ATX_FLAG_PDL) ? 1 : 0;
    const int kmode = mode | (ordered ? 0x100 : 0);
    if (T <= 16) return launch_gemm<16>(p, 0, out, ldo, T, kmode, splits, pdl, stream);
    if (T <= 32) return launch_gemm<32>(p, 1, out, ldo, T, kmode, splits, pdl, stream);
    if (T <= 64) return launch_gemm<64>(p, 2, out, ldo, T, kmode, splits, pdl, stream);
    // prefill-sized batches: the persistent, warp-specialised kernel (one CTA per SM, two TMEM accumulators)
    ATX_REQUIRE(splits == 1, "atx_gemm_run: split-K is for decode-sized steps (T <= 64); got T=%d splits=%d", T, splits);
    if (T <= 128) return launch_gemm_persistent<128>(p, 3, out, ldo, T, mode, pdl, stream);
    return launch_gemm_persistent<256>(p, 4, out, ldo, T, mode, pdl, stream);
}

}  // extern "C"
