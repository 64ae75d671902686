// Tensor-parallel all-reduce fused with the residual add and the RMSNorm that consumes it.
//
// In the reference's stack (vLLM behind llm/serve_llm.py:374-379) a tensor-parallel Llama layer does
// o_proj / down_proj -> NCCL all-reduce -> residual add -> RMSNorm (transformers modeling_llama.py:313-332
// for the add + norm). At decode sizes the message is tiny ([tokens, hidden] fp32: 1 MiB at 32 tokens of a
// 70B model) and 160 of them per step are pure latency. Here the collective, the add and the norm are ONE
// kernel over NVLink peer memory:
//
//   every rank's GEMM accumulates its partial projection into its own symmetric buffer `part`;
//   this kernel 1) tells every peer "my partial is complete" and waits for the same from all of them
//   (release/acquire flags in peer memory), 2) each CTA pulls one token row from ALL ranks' buffers with
//   16-byte peer loads, adds them to the local fp32 residual row, 3) writes the residual back and the
//   bf16 RMS-normalised row for the next GEMM, 4) clears the OTHER of the two ping-pong partial buffers
//   (every peer finished reading it: they could not have signalled this round otherwise).
//
// One-shot (every rank pulls everything) is latency-optimal for the <= 256-token steps this is used for;
// prefill-sized steps keep NCCL's bandwidth-optimal ring. Ranks must be one process per GPU: the flag
// wait spins on values that only the peers' kernels write.
#include "common.cuh"

#include "../../include/atx_b200.h"

namespace atx {

constexpr int kArThreads = 512;
constexpr int kArMaxVec = 4;   // float4 per thread -> H <= 8192
constexpr int kArMaxRanks = 8;

struct ArArgs {
    const float* part[kArMaxRanks];   // every rank's active partial buffer (peer-mapped), [T, H]
    int* flags[kArMaxRanks];          // every rank's flag array (peer-mapped), int[world]
};

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

__global__ void __launch_bounds__(kArThreads)
allreduce_add_rmsnorm_kernel(ArArgs a, float* __restrict__ x, float* __restrict__ clear_buf, int clear_rows,
                             int* __restrict__ state, int rank, int world, const int* __restrict__ rows,
                             const uint2* __restrict__ w, uint2* __restrict__ out, int T, int H, float eps,
                             uint4* __restrict__ zero_ptr, long long zero_vecs) {
    __shared__ float red[kArThreads / 32];
    pdl_launch_dependents();
    pdl_wait();  // this rank's own GEMM has finished adding into its partial buffer
    const int epoch = *reinterpret_cast<volatile int*>(state) + 1;  // same in every CTA: bumped by the last CTA out
    if (blockIdx.x >= T) {
        const long long nb = gridDim.x - T;
        for (long long i = static_cast<long long>(blockIdx.x - T) * blockDim.x + threadIdx.x; i < zero_vecs;
             i += nb * blockDim.x)
            zero_ptr[i] = make_uint4(0, 0, 0, 0);
    } else {
        // ---- cross-GPU barrier: announce, then wait for every rank's announcement of this round
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

        const int t = blockIdx.x;
        const int r = rows ? rows[t] : t;
        const int nv = H / 4;
        float4* xr = reinterpret_cast<float4*>(x + static_cast<long long>(r) * H);
        float4 v[kArMaxVec];
        float ss = 0.f;
#pragma unroll
        for (int j = 0; j < kArMaxVec; ++j) {
            const int i = threadIdx.x + j * kArThreads;
            if (i < nv) {
                float4 acc = xr[i];
#pragma unroll
                for (int p = 0; p < kArMaxRanks; ++p) {
                    if (p < world) {
                        const float4 q = ld_peer_f4(a.part[p] + static_cast<long long>(r) * H + 4 * i);
                        acc.x += q.x; acc.y += q.y; acc.z += q.z; acc.w += q.w;
                    }
                }
                v[j] = acc;
                if (!rows) xr[i] = acc;  // the residual stream now includes the reduced projection
                ss += acc.x * acc.x + acc.y * acc.y + acc.z * acc.z + acc.w * acc.w;
            }
        }
        ss = warp_sum(ss);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
        __syncthreads();
        float tot = 0.f;
#pragma unroll
        for (int i = 0; i < kArThreads / 32; ++i) tot += red[i];
        const float rs = rsqrtf(tot / static_cast<float>(H) + eps);
        uint2* orow = out + static_cast<long long>(t) * nv;
#pragma unroll
        for (int j = 0; j < kArMaxVec; ++j) {
            const int i = threadIdx.x + j * kArThreads;
            if (i < nv) {
                const uint2 wv = __ldg(w + i);
                uint2 o;
                o.x = pack_bf16(bf16lo(wv.x) * (v[j].x * rs), bf16hi(wv.x) * (v[j].y * rs));
                o.y = pack_bf16(bf16lo(wv.y) * (v[j].z * rs), bf16hi(wv.y) * (v[j].w * rs));
                orow[i] = o;
            }
        }
        // the other ping-pong buffer was fully read by every peer in the previous round: hand it back clean
        if (clear_buf) {
            for (int rr = blockIdx.x; rr < clear_rows; rr += T) {
                float4* c = reinterpret_cast<float4*>(clear_buf + static_cast<long long>(rr) * H);
                for (int i = threadIdx.x; i < nv; i += kArThreads) c[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
    }
    // ---- last CTA out publishes the round number for the next call
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const int prev = atomicAdd(state + 1, 1);
        if (prev == static_cast<int>(gridDim.x) - 1) {
            state[1] = 0;
            __threadfence();
            *reinterpret_cast<volatile int*>(state) = epoch;
        }
    }
}

}  // namespace atx

using namespace atx;

extern "C" {

int atx_allreduce_add_rmsnorm(float* x, const void* const* part_ptrs, void* const* flag_ptrs, void* clear_buf,
                              int clear_rows, int* state, int rank, int world, const int* rows, const void* w,
                              void* out, int T, int H, float eps, void* zero_ptr, long long zero_bytes, int flags,
                              atx_stream_t stream) {
    ATX_REQUIRE(x && part_ptrs && flag_ptrs && state && w && out, "atx_allreduce_add_rmsnorm: null pointer");
    ATX_REQUIRE(world >= 2 && world <= kArMaxRanks && rank >= 0 && rank < world,
                "atx_allreduce_add_rmsnorm: world=%d (2..%d), rank=%d", world, kArMaxRanks, rank);
    ATX_REQUIRE(T > 0 && H > 0 && H % 4 == 0 && H <= kArThreads * kArMaxVec * 4,
                "atx_allreduce_add_rmsnorm: bad shape T=%d H=%d (H %% 4 == 0, H <= %d)", T, H, kArThreads * kArMaxVec * 4);
    ATX_REQUIRE(zero_bytes >= 0 && zero_bytes % 16 == 0 && (zero_bytes == 0 || zero_ptr),
                "atx_allreduce_add_rmsnorm: zero_bytes=%lld must be a non-negative multiple of 16 with a buffer",
                zero_bytes);
    ArArgs a = {};
    for (int p = 0; p < world; ++p) {
        ATX_REQUIRE(part_ptrs[p] && flag_ptrs[p], "atx_allreduce_add_rmsnorm: null peer pointer for rank %d", p);
        a.part[p] = static_cast<const float*>(part_ptrs[p]);
        a.flags[p] = static_cast<int*>(flag_ptrs[p]);
    }
    const long long zv = zero_bytes / 16;
    long long zb = (zv + kArThreads * 8 - 1) / (kArThreads * 8);
    if (zb > 1024) zb = 1024;
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = (flags & ATX_FLAG_PDL) ? 1 : 0;
    cfg.gridDim = dim3(static_cast<unsigned>(T + zb));
    cfg.blockDim = dim3(kArThreads);
    cfg.stream = reinterpret_cast<cudaStream_t>(stream);
    ATX_CUDA(cudaLaunchKernelEx(&cfg, allreduce_add_rmsnorm_kernel, a, x, static_cast<float*>(clear_buf), clear_rows,
                                state, rank, world, rows, reinterpret_cast<const uint2*>(w),
                                reinterpret_cast<uint2*>(out), T, H, eps, reinterpret_cast<uint4*>(zero_ptr), zv));
    return ATX_OK;
}

}  // extern "C"
