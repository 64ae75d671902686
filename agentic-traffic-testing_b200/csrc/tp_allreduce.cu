// Tensor-parallel all-reduce fused with the residual add and the RMSNorm that consumes it.
//
// In the reference's stack (vLLM behind llm/serve_llm.py:374-379) a tensor-parallel Llama layer does
// o_proj / down_proj -> NCCL all-reduce -> residual add -> RMSNorm (transformers modeling_llama.py:313-332
// for the add + norm). At decode sizes the message is tiny ([tokens, hidden] fp32: 1 MiB at 32 tokens of a
// 70B model) and 160 of them per step are pure latency. Here the collective, the add and the norm are ONE
// kernel over NVLink peer memory, in two forms (both two-shot: rank r owns the token rows t = r mod world):
//   * <= 64 rows: the low-latency push form (atx_allreduce_add_rmsnorm_ll): no flag exchange at all, the data
//     is the flag (sentinel-filled receive slots), one NVLink one-way latency per phase: 11 us per call;
//   * <= 256 rows: the flag-barrier form (atx_reduce_scatter_rmsnorm_all_gather): release/acquire flags in peer
//     memory, pulls of the owned rows, peer stores of the normalised bf16 rows: 22 us per call;
// prefill-sized steps keep NCCL's bandwidth-optimal ring. A one-shot form (every rank pulls everything, one
// cluster per row) was measured at 12 us per call and 8.9 ms per 70B step against 6.3 with the push form and is
// gone (profiles/r01k, r01m). Ranks must be one process per GPU: the waits spin on values that only the peers'
// kernels write.
#include "common.cuh"

#include "../../include/atx_b200.h"

namespace atx {

constexpr int kArThreads = 512;
constexpr int kArMaxCluster = 8;   // CTAs per token row, one float4 per thread -> H <= 16384
constexpr int kArMaxRanks = 8;

__device__ __forceinline__ void st_release_sys(int* p, int v) {
    asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
    int v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 ld_peer_f4(const float* p) {
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(p)
                 : "memory");
    return v;
}

__device__ __forceinline__ void ar_cluster_barrier() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ float ar_dsmem_ld(const float* local_ptr, uint32_t cta_rank) {
    uint32_t remote;
    float v;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(local_ptr)), "r"(cta_rank));
    asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(remote) : "memory");
    return v;
}

// ---- two-shot variant: reduce-scatter by token row, RMSNorm on the owner, all-gather of the bf16 result ----
// One-shot makes every rank pull the whole [T, H] fp32 message from every peer ((world-1) x T x H x 4 bytes
// in, 7 MiB at 32 tokens of a 70B model on 8 GPUs). Here rank r owns the token rows t = r (mod world): it pulls
// only those rows (1/world of the message), keeps the fp32 residual of ITS rows, and pushes the normalised
// bf16 rows straight into every rank's activation buffer (the next GEMM's TMA operand) with peer stores:
// (world-1)/world x T x H x (4 + 2) bytes per rank, 8x less NVLink traffic at world 8. A second cross-GPU
// barrier, run by the last CTA out, holds the grid open until every peer's rows have landed here, so the
// consumer needs nothing but stream order. The residual rows a rank does not own go stale and are never read.
struct Ar2Args {
    const float* part[kArMaxRanks];
    int* flags[kArMaxRanks];          // int[2 * world]: [0, world) barrier 1, [world, 2 world) barrier 2
    uint2* out[kArMaxRanks];          // every rank's activation buffer (peer-mapped), bf16 [T, H]
};

__device__ __forceinline__ void st_peer_u2(uint2* p, uint2 v) {
    asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}

__global__ void __launch_bounds__(kArThreads)
reduce_scatter_rmsnorm_all_gather_kernel(Ar2Args a, float* __restrict__ x, float* __restrict__ clear_buf,
                                         int clear_rows, int* __restrict__ state, int rank, int world,
                                         const int* __restrict__ rows, const uint2* __restrict__ w, int T, int H,
                                         float eps, uint4* __restrict__ zero_ptr, long long zero_vecs, int C,
                                         int row_clusters) {
    __shared__ float red[kArThreads / 32];
    __shared__ float css;
    __shared__ int is_last;
    pdl_launch_dependents();
    pdl_wait();
    const int epoch = *reinterpret_cast<volatile int*>(state) + 1;
    const int j = blockIdx.x / C, c = blockIdx.x % C;
    if (j >= row_clusters) {
        const long long first = static_cast<long long>(row_clusters) * C;
        const long long nb = gridDim.x - first;
        for (long long i = (blockIdx.x - first) * blockDim.x + threadIdx.x; i < zero_vecs; i += nb * blockDim.x)
            zero_ptr[i] = make_uint4(0, 0, 0, 0);
    } else {
        // ---- barrier 1: every rank's GEMM has finished adding into its partial buffer
        if (blockIdx.x == 0 && threadIdx.x < world) {
            __threadfence_system();
            st_release_sys(a.flags[threadIdx.x] + rank, epoch);
        }
        if (threadIdx.x < world) {
            const int* f = a.flags[rank] + threadIdx.x;
            while (ld_acquire_sys(f) < epoch) {
            }
        }
        __syncthreads();
        // the output row this cluster produces (uniform over the cluster). Ownership follows the RESIDUAL
        // row: without a gather rank r owns t = r (mod world); with one, cluster j looks at output row j and
        // only the rank that owns residual row rows[j] (whose x is current there) produces it.
        const int t = rows ? j : rank + j * world;
        const int r = (t < T) ? (rows ? rows[t] : t) : 0;
        if (t < T && (!rows || r % world == rank)) {
            const int nv = H / 4, per = (nv + C - 1) / C;
            const int i = c * per + threadIdx.x;
            const bool ok = threadIdx.x < per && i < nv;
            float4* xr = reinterpret_cast<float4*>(x + static_cast<long long>(r) * H);
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            float ss = 0.f;
            if (ok) {
                float4 q[kArMaxRanks];
#pragma unroll
                for (int p = 0; p < kArMaxRanks; ++p)
                    if (p < world) q[p] = ld_peer_f4(a.part[p] + static_cast<long long>(r) * H + 4 * i);
                acc = xr[i];
#pragma unroll
                for (int p = 0; p < kArMaxRanks; ++p) {   // fixed rank order
                    if (p < world) {
                        acc.x += q[p].x; acc.y += q[p].y; acc.z += q[p].z; acc.w += q[p].w;
                    }
                }
                if (!rows) xr[i] = acc;
                ss = acc.x * acc.x + acc.y * acc.y + acc.z * acc.z + acc.w * acc.w;
            }
            ss = warp_sum(ss);
            if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
            __syncthreads();
            if (threadIdx.x == 0) {
                float tot = 0.f;
#pragma unroll
                for (int k = 0; k < kArThreads / 32; ++k) tot += red[k];
                css = tot;
            }
            float tot;
            if (C > 1) {
                ar_cluster_barrier();
                tot = 0.f;
                for (int k = 0; k < C; ++k) tot += ar_dsmem_ld(&css, static_cast<uint32_t>(k));
            } else {
                __syncthreads();
                tot = css;
            }
            const float rs = rsqrtf(tot / static_cast<float>(H) + eps);
            if (ok) {
                const uint2 wv = __ldg(w + i);
                uint2 o;
                o.x = pack_bf16(bf16lo(wv.x) * (acc.x * rs), bf16hi(wv.x) * (acc.y * rs));
                o.y = pack_bf16(bf16lo(wv.y) * (acc.z * rs), bf16hi(wv.y) * (acc.w * rs));
                const long long at = static_cast<long long>(t) * nv + i;
#pragma unroll
                for (int p = 0; p < kArMaxRanks; ++p)
                    if (p < world) st_peer_u2(a.out[p] + at, o);   // all-gather by peer stores
            }
            if (C > 1) ar_cluster_barrier();   // peers may still be reading css
        }
        if (clear_buf) {
            const long long cv = static_cast<long long>(clear_rows) * (H / 4);
            float4* cb = reinterpret_cast<float4*>(clear_buf);
            for (long long k = static_cast<long long>(blockIdx.x) * kArThreads + threadIdx.x; k < cv;
                 k += static_cast<long long>(row_clusters) * C * kArThreads)
                cb[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    // ---- last CTA out: barrier 2 (every rank's rows have been pushed everywhere), then publish the round
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();   // this CTA's peer stores (ordered before by the barrier above) become visible
        const int prev = atomicAdd(state + 1, 1);
        is_last = prev == static_cast<int>(gridDim.x) - 1;
    }
    __syncthreads();
    if (is_last) {
        if (threadIdx.x < world) {
            __threadfence_system();
            st_release_sys(a.flags[threadIdx.x] + world + rank, epoch);
            const int* f = a.flags[rank] + world + threadIdx.x;
            while (ld_acquire_sys(f) < epoch) {
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            state[1] = 0;
            __threadfence();
            *reinterpret_cast<volatile int*>(state) = epoch;
        }
    }
}

// ---- low-latency variant: push + "the data is the flag" (no cross-GPU flag barrier, no pulls) -------------
// The two kernels above pay a flag exchange (fence + store + acquire spin) and a pull round trip per call:
// 12 us (one-shot) / 22 us (two-shot) measured back to back on B200 for a 1 MiB message, more than the GEMM
// they follow. Here every buffer a peer writes into is kept filled with a SENTINEL bit pattern (-0.0, which a
// sanitised payload never contains) and the receiver simply spins on the data:
//   A  every rank reads its partial rows (local) and pushes row t to the rank that owns it, into that rank's
//      receive slot [source][t] (fire-and-forget peer stores);
//   B  the owner spins until all `world` slots of its rows are non-sentinel, re-arms them, adds them to its
//      fp32 residual row in fixed source order, normalises, and pushes the bf16 row into EVERY rank's
//      activation buffer;
//   C  every rank spins until all T rows of its activation buffer have arrived, then the grid ends and the
//      next GEMM (plain stream order) reads them with TMA.
// One NVLink one-way latency per phase instead of a fence + flag + round trip. Re-arming: receive slots are
// reset by their reader right after reading (two parities alternate); the activation buffer of the OTHER
// parity (consumed by the GEMMs that ran before this call) is reset here via reset_ptr. A peer can only be
// one call ahead of this rank (it needs this rank's phase-A rows to finish a call), so two parities suffice.
constexpr uint32_t kSentinelF32 = 0x80000000u;
constexpr uint32_t kSentinelBf16x2 = 0x80008000u;

struct LlArgs {
    float* rs[kArMaxRanks];    // rank p's receive slots of this parity (peer-mapped): fp32 [world][T_cap][H]
    uint2* out[kArMaxRanks];   // rank p's activation buffer of this parity (peer-mapped): bf16 [>= T, H]
};

__device__ __forceinline__ void st_peer_f4(float* p, float4 v) {
    asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}
__device__ __forceinline__ uint2 ld_poll_u2(const uint2* p) {
    uint2 v;
    asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float unsentinel(float v) {   // -0.0 -> +0.0, everything else unchanged
    return __float_as_uint(v) == kSentinelF32 ? 0.f : v;
}
__device__ __forceinline__ uint32_t unsentinel_bf16x2(uint32_t u) {
    if ((u & 0xffffu) == 0x8000u) u &= 0xffff0000u;
    if ((u >> 16) == 0x8000u) u &= 0x0000ffffu;
    return u;
}
__device__ __forceinline__ bool arrived(float4 q) {
    return __float_as_uint(q.x) != kSentinelF32 && __float_as_uint(q.y) != kSentinelF32 &&
           __float_as_uint(q.z) != kSentinelF32 && __float_as_uint(q.w) != kSentinelF32;
}
__device__ __forceinline__ bool arrived(uint2 u) {
    return (u.x & 0xffffu) != 0x8000u && (u.x >> 16) != 0x8000u && (u.y & 0xffffu) != 0x8000u && (u.y >> 16) != 0x8000u;
}

__global__ void __launch_bounds__(kArThreads)
allreduce_ll_kernel(LlArgs a, float* __restrict__ x, const float* __restrict__ part, float* __restrict__ clear_buf,
                    int clear_rows, uint4* __restrict__ reset_ptr, long long reset_vecs, int T_cap, int rank, int world,
                    const int* __restrict__ rows, const uint2* __restrict__ w, int T, int H, float eps,
                    uint4* __restrict__ zero_ptr, long long zero_vecs, int C) {
    __shared__ float red[kArThreads / 32];
    __shared__ float css;
    pdl_launch_dependents();
    pdl_wait();   // this rank's GEMM has finished adding into `part`; the GEMMs reading the other parity are done
    const int t = blockIdx.x / C, c = blockIdx.x % C;
    const int nv = H / 4;
    const long long gtid = static_cast<long long>(blockIdx.x) * kArThreads + threadIdx.x;
    const long long gthreads = static_cast<long long>(gridDim.x) * kArThreads;
    if (t < T) {
        const int r = rows ? rows[t] : t;
        const int owner = r % world;           // ownership follows the residual row (uniform over the cluster)
        const int per = (nv + C - 1) / C;
        const int i = c * per + threadIdx.x;
        const bool ok = threadIdx.x < per && i < nv;
        // ---- A: my partial of row t goes to its owner
        if (ok) {
            float4 v = *reinterpret_cast<const float4*>(part + static_cast<long long>(r) * H + 4 * i);
            v.x = unsentinel(v.x); v.y = unsentinel(v.y); v.z = unsentinel(v.z); v.w = unsentinel(v.w);
            st_peer_f4(a.rs[owner] + (static_cast<long long>(rank) * T_cap + t) * H + 4 * i, v);
        }
    }
    // ---- housekeeping while the rows travel: re-arm the other parity's activation buffer, clear the other
    // partial buffer, zero the split-K target of the next GEMM (all local)
    {
        const uint4 sent = make_uint4(kSentinelBf16x2, kSentinelBf16x2, kSentinelBf16x2, kSentinelBf16x2);
        for (long long k = gtid; k < reset_vecs; k += gthreads) reset_ptr[k] = sent;
        if (clear_buf) {
            const long long cv = static_cast<long long>(clear_rows) * nv;
            float4* cb = reinterpret_cast<float4*>(clear_buf);
            for (long long k = gtid; k < cv; k += gthreads) cb[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        for (long long k = gtid; k < zero_vecs; k += gthreads) zero_ptr[k] = make_uint4(0, 0, 0, 0);
    }
    if (t >= T) return;
    {
        const int r = rows ? rows[t] : t;
        const int owner = r % world;
        const int per = (nv + C - 1) / C;
        const int i = c * per + threadIdx.x;
        const bool ok = threadIdx.x < per && i < nv;
        // ---- B: the owner waits for all sources, reduces, normalises, broadcasts
        if (owner == rank) {
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            float ss = 0.f;
            if (ok) {
                float4* xr = reinterpret_cast<float4*>(x + static_cast<long long>(r) * H);
                acc = xr[i];
                float* mine = a.rs[rank] + static_cast<long long>(t) * H + 4 * i;
                const long long src_stride = static_cast<long long>(T_cap) * H;
                float4 q[kArMaxRanks];
                uint32_t pending = (1u << world) - 1u;
                while (pending) {
#pragma unroll
                    for (int p = 0; p < kArMaxRanks; ++p) {
                        if (p < world && (pending >> p & 1u)) {
                            q[p] = ld_peer_f4(mine + p * src_stride);
                            if (arrived(q[p])) pending &= ~(1u << p);
                        }
                    }
                }
                const float4 sent = make_float4(__uint_as_float(kSentinelF32), __uint_as_float(kSentinelF32),
                                                __uint_as_float(kSentinelF32), __uint_as_float(kSentinelF32));
#pragma unroll
                for (int p = 0; p < kArMaxRanks; ++p) {   // fixed source order: every run sums identically
                    if (p < world) {
                        *reinterpret_cast<float4*>(mine + p * src_stride) = sent;   // re-arm for the call after next
                        acc.x += q[p].x; acc.y += q[p].y; acc.z += q[p].z; acc.w += q[p].w;
                    }
                }
                if (!rows) xr[i] = acc;
                ss = acc.x * acc.x + acc.y * acc.y + acc.z * acc.z + acc.w * acc.w;
            }
            ss = warp_sum(ss);
            if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
            __syncthreads();
            if (threadIdx.x == 0) {
                float tot = 0.f;
#pragma unroll
                for (int k = 0; k < kArThreads / 32; ++k) tot += red[k];
                css = tot;
            }
            float tot;
            if (C > 1) {
                ar_cluster_barrier();
                tot = 0.f;
                for (int k = 0; k < C; ++k) tot += ar_dsmem_ld(&css, static_cast<uint32_t>(k));
            } else {
                __syncthreads();
                tot = css;
            }
            const float rs = rsqrtf(tot / static_cast<float>(H) + eps);
            if (ok) {
                const uint2 wv = __ldg(w + i);
                uint2 o;
                o.x = unsentinel_bf16x2(pack_bf16(bf16lo(wv.x) * (acc.x * rs), bf16hi(wv.x) * (acc.y * rs)));
                o.y = unsentinel_bf16x2(pack_bf16(bf16lo(wv.y) * (acc.z * rs), bf16hi(wv.y) * (acc.w * rs)));
                const long long at = static_cast<long long>(t) * nv + i;
#pragma unroll
                for (int p = 0; p < kArMaxRanks; ++p)
                    if (p < world) st_peer_u2(a.out[p] + at, o);
            }
            if (C > 1) ar_cluster_barrier();   // peers may still be reading css
        }
        // ---- C: this rank's copy of row t has arrived (from its owner, possibly this rank itself)
        if (ok) {
            const uint2* mine = a.out[rank] + static_cast<long long>(t) * nv + i;
            while (!arrived(ld_poll_u2(mine))) {
            }
        }
    }
}

}  // namespace atx

using namespace atx;

extern "C" {

int atx_reduce_scatter_rmsnorm_all_gather(float* x, const void* const* part_ptrs, void* const* flag_ptrs,
                                          void* const* out_ptrs, void* clear_buf, int clear_rows, int* state, int rank,
                                          int world, const int* rows, const void* w, int T, int H, float eps,
                                          void* zero_ptr, long long zero_bytes, int flags, atx_stream_t stream) {
    ATX_REQUIRE(x && part_ptrs && flag_ptrs && out_ptrs && state && w,
                "atx_reduce_scatter_rmsnorm_all_gather: null pointer");
    ATX_REQUIRE(world >= 2 && world <= kArMaxRanks && rank >= 0 && rank < world,
                "atx_reduce_scatter_rmsnorm_all_gather: world=%d (2..%d), rank=%d", world, kArMaxRanks, rank);
    ATX_REQUIRE(T > 0 && H > 0 && H % 4 == 0 && H <= kArThreads * kArMaxCluster * 4,
                "atx_reduce_scatter_rmsnorm_all_gather: bad shape T=%d H=%d (H %% 4 == 0, H <= %d)", T, H,
                kArThreads * kArMaxCluster * 4);
    ATX_REQUIRE(zero_bytes >= 0 && zero_bytes % 16 == 0 && (zero_bytes == 0 || zero_ptr),
                "atx_reduce_scatter_rmsnorm_all_gather: zero_bytes=%lld must be a non-negative multiple of 16 with a "
                "buffer", zero_bytes);
    Ar2Args a = {};
    for (int p = 0; p < world; ++p) {
        ATX_REQUIRE(part_ptrs[p] && flag_ptrs[p] && out_ptrs[p],
                    "atx_reduce_scatter_rmsnorm_all_gather: null peer pointer for rank %d", p);
        a.part[p] = static_cast<const float*>(part_ptrs[p]);
        a.flags[p] = static_cast<int*>(flag_ptrs[p]);
        a.out[p] = static_cast<uint2*>(out_ptrs[p]);
    }
    const long long zv = zero_bytes / 16;
    long long zb = (zv + kArThreads * 8 - 1) / (kArThreads * 8);
    if (zb > 1024) zb = 1024;
    int C = 1;
    while (C * kArThreads * 4 < H) C *= 2;
    const int row_clusters = rows ? T : (T + world - 1) / world;   // rows t = rank, rank + world, ... (or all, filtered)
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[2];
    int na = 0;
    if (flags & ATX_FLAG_PDL) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    if (C > 1) {
        attr[na].id = cudaLaunchAttributeClusterDimension;
        attr[na].val.clusterDim.x = static_cast<unsigned>(C);
        attr[na].val.clusterDim.y = 1;
        attr[na].val.clusterDim.z = 1;
        ++na;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    zb = (zb + C - 1) / C;
    cfg.gridDim = dim3(static_cast<unsigned>((row_clusters + zb) * C));
    cfg.blockDim = dim3(kArThreads);
    cfg.stream = reinterpret_cast<cudaStream_t>(stream);
    ATX_CUDA(cudaLaunchKernelEx(&cfg, reduce_scatter_rmsnorm_all_gather_kernel, a, x, static_cast<float*>(clear_buf),
                                clear_rows, state, rank, world, rows, reinterpret_cast<const uint2*>(w), T, H, eps,
                                reinterpret_cast<uint4*>(zero_ptr), zv, C, row_clusters));
    return ATX_OK;
}

int atx_allreduce_add_rmsnorm_ll(float* x, const float* part, void* const* rs_ptrs, void* const* out_ptrs,
                                 void* clear_buf, int clear_rows, void* reset_ptr, long long reset_bytes, int T_cap,
                                 int rank, int world, const int* rows, const void* w, int T, int H, float eps,
                                 void* zero_ptr, long long zero_bytes, int flags, atx_stream_t stream) {
    ATX_REQUIRE(x && part && rs_ptrs && out_ptrs && w, "atx_allreduce_add_rmsnorm_ll: null pointer");
    ATX_REQUIRE(world >= 2 && world <= kArMaxRanks && rank >= 0 && rank < world,
                "atx_allreduce_add_rmsnorm_ll: world=%d (2..%d), rank=%d", world, kArMaxRanks, rank);
    ATX_REQUIRE(T > 0 && T <= T_cap && H > 0 && H % 4 == 0 && H <= kArThreads * kArMaxCluster * 4,
                "atx_allreduce_add_rmsnorm_ll: bad shape T=%d (cap %d) H=%d (H %% 4 == 0, H <= %d)", T, T_cap, H,
                kArThreads * kArMaxCluster * 4);
    ATX_REQUIRE(zero_bytes >= 0 && zero_bytes % 16 == 0 && (zero_bytes == 0 || zero_ptr) && reset_bytes >= 0 &&
                    reset_bytes % 16 == 0 && (reset_bytes == 0 || reset_ptr),
                "atx_allreduce_add_rmsnorm_ll: zero_bytes=%lld / reset_bytes=%lld must be non-negative multiples of 16 "
                "with a buffer", zero_bytes, reset_bytes);
    int C = 1;
    while (C * kArThreads * 4 < H) C *= 2;
    // every CTA of the grid spins on data that other GPUs produce: the whole grid must be co-resident
    ATX_REQUIRE(static_cast<long long>(T) * C <= 2ll * num_sms(),
                "atx_allreduce_add_rmsnorm_ll: T=%d rows x %d CTAs exceed the co-resident grid (%d)", T, C, 2 * num_sms());
    LlArgs a = {};
    for (int p = 0; p < world; ++p) {
        ATX_REQUIRE(rs_ptrs[p] && out_ptrs[p], "atx_allreduce_add_rmsnorm_ll: null peer pointer for rank %d", p);
        a.rs[p] = static_cast<float*>(rs_ptrs[p]);
        a.out[p] = static_cast<uint2*>(out_ptrs[p]);
    }
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[2];
    int na = 0;
    if (flags & ATX_FLAG_PDL) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    if (C > 1) {
        attr[na].id = cudaLaunchAttributeClusterDimension;
        attr[na].val.clusterDim.x = static_cast<unsigned>(C);
        attr[na].val.clusterDim.y = 1;
        attr[na].val.clusterDim.z = 1;
        ++na;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    // a few extra clusters so the housekeeping stores are spread even when T is 1
    const int extra = T * C < 32 ? (32 - T * C + C - 1) / C : 0;
    cfg.gridDim = dim3(static_cast<unsigned>((T + extra) * C));
    cfg.blockDim = dim3(kArThreads);
    cfg.stream = reinterpret_cast<cudaStream_t>(stream);
    ATX_CUDA(cudaLaunchKernelEx(&cfg, allreduce_ll_kernel, a, x, part, static_cast<float*>(clear_buf), clear_rows,
                                reinterpret_cast<uint4*>(reset_ptr), reset_bytes / 16, T_cap, rank, world, rows,
                                reinterpret_cast<const uint2*>(w), T, H, eps, reinterpret_cast<uint4*>(zero_ptr),
                                zero_bytes / 16, C));
    return ATX_OK;
}

}  // extern "C"
