// Token sampler over fp32 logits: greedy arg-max, temperature sampling, top-p (nucleus) sampling.
//
// Replaces vLLM's sampler for SamplingParams(temperature=0.2) (llm/serve_llm.py:379, 521-525; top_p
// defaults to 1) and HF's do_sample / temperature=0.7 (llm/hf_cpu_server.py:91); arg-max is the
// greedy rule the parity tests use (torch.argmax: lowest index wins ties).
//
// A categorical draw from softmax(l / T) is made with the Gumbel-max rule, arg-max_i(l_i / T + g_i),
// g_i = -log(-log(u_i)), which needs no normalisation, no prefix sums and no host round trip. u_i
// comes from a counter-based integer hash of (seed, step, i) restated in oracle/sampler_ref.py.
// The nucleus {i : sum_{p_j > p_i} p_j < top_p} (the HF / vLLM top-p rule) is found by bisection on
// the BIT PATTERN of the un-normalised probability (positive floats order like their bits), i.e. an
// exact threshold in at most 31 passes over logits that stay L2-resident.
#include "common.cuh"

#include "../../include/atx_b200.h"

namespace atx {

constexpr int kSampThreads = 1024;

__device__ __forceinline__ uint32_t lowbias32(uint32_t x) {
    x ^= x >> 16;
    x *= 0x7feb352du;
    x ^= x >> 15;
    x *= 0x846ca68bu;
    x ^= x >> 16;
    return x;
}
__device__ __forceinline__ float gumbel(uint32_t key, uint32_t idx) {
    const uint32_t x = lowbias32(idx ^ key);
    const float u = (static_cast<float>(x >> 8) + 0.5f) * 5.9604644775390625e-08f;  // (0, 1)
    return -logf(-logf(u));
}

struct ArgMax {
    float v;
    int i;
};
__device__ __forceinline__ ArgMax better(ArgMax a, ArgMax b) {
    return (b.v > a.v || (b.v == a.v && b.i < a.i)) ? b : a;
}
__device__ __forceinline__ ArgMax block_argmax(ArgMax a, ArgMax* sh) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        ArgMax b;
        b.v = __shfl_xor_sync(0xffffffffu, a.v, o);
        b.i = __shfl_xor_sync(0xffffffffu, a.i, o);
        a = better(a, b);
    }
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = a;
    __syncthreads();
    ArgMax r = sh[0];
    for (int w = 1; w < kSampThreads / 32; ++w) r = better(r, sh[w]);
    return r;
}
__device__ __forceinline__ float block_sum(float v, float* sh) {
    v = warp_sum(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    float r = 0.f;
    for (int w = 0; w < kSampThreads / 32; ++w) r += sh[w];
    return r;
}

// Visit every logit of a row: 16-byte loads, four in flight per thread, when the row allows it (the 128 256-wide
// rows of the Llama vocabulary do). A pass over a row is a chain of L2 round trips, not bandwidth: with scalar
// loads one after another the arg-max pass alone took 20 us.
template <class F>
__device__ __forceinline__ void for_each_logit(const float* __restrict__ row, int V, bool vec, F f) {
    if (vec) {
        const float4* r4 = reinterpret_cast<const float4*>(row);
        const int n4 = V >> 2;
#pragma unroll 4
        for (int j = threadIdx.x; j < n4; j += kSampThreads) {
            const float4 v = __ldg(r4 + j);
            f(v.x, 4 * j);
            f(v.y, 4 * j + 1);
            f(v.z, 4 * j + 2);
            f(v.w, 4 * j + 3);
        }
    } else {
        for (int i = threadIdx.x; i < V; i += kSampThreads) f(row[i], i);
    }
}

// With `pairs` set the row is one SHARD of a vocabulary-parallel row (columns [idx_offset, idx_offset + V) of the
// full row): instead of a token id the kernel emits the shard's best (score, global index) pair, score = the
// logit (greedy) or logit / T + Gumbel noise keyed by the GLOBAL index; the arg-max over the shards' pairs
// (sample_merge_kernel) is then exactly the token the unsharded kernel draws. No nucleus here (it needs the sum over
// the whole row): the caller assembles full rows when a request asks for top_p < 1.
__global__ void __launch_bounds__(kSampThreads)
sample_kernel(const float* __restrict__ logits, long long ldl, const float* __restrict__ temperature,
              const float* __restrict__ top_p, const unsigned long long* __restrict__ seeds,
              const int* __restrict__ steps, int* __restrict__ out_ids, int V, int idx_offset,
              float2* __restrict__ pairs) {
    __shared__ ArgMax sh_am[kSampThreads / 32];
    __shared__ float sh_f[kSampThreads / 32];
    pdl_launch_dependents();
    pdl_wait();
    const int b = blockIdx.x;
    const float* row = logits + static_cast<long long>(b) * ldl;
    const float T = temperature ? temperature[b] : 0.f;
    const bool vec = (V & 3) == 0 && (reinterpret_cast<uintptr_t>(row) & 15) == 0;

    // pass 1: arg-max of the raw logits (also the softmax shift)
    ArgMax am = {-INFINITY, 0x7fffffff};
    for_each_logit(row, V, vec, [&](float l, int i) { am = better(am, ArgMax{l, i}); });
    am = block_argmax(am, sh_am);
    if (!(T > 0.f)) {
        if (threadIdx.x == 0) {
            if (pairs) pairs[b] = make_float2(am.v, __int_as_float(am.i + idx_offset));
            else out_ids[b] = am.i;
        }
        return;
    }
    const float inv_t = 1.f / T;
    const float lmax = am.v;
    const float P = (top_p && !pairs) ? top_p[b] : 1.f;

    float thr = 0.f;  // keep e_i >= thr, e_i = exp((l_i - lmax) / T) in (0, 1]
    if (P < 1.f) {
        float z = 0.f;
        for_each_logit(row, V, vec, [&](float l, int) { z += expf((l - lmax) * inv_t); });
        z = block_sum(z, sh_f);
        const float target = P * z;
        // smallest float x in [0, 1] with sum_{e_j > x} e_j < target
        uint32_t lo = 0u, hi = __float_as_uint(1.0f);  // f(lo) false unless target > z; f(hi) true (empty sum)
        while (hi - lo > 1u) {
            const uint32_t mid = lo + (hi - lo) / 2;
            const float x = __uint_as_float(mid);
            float s = 0.f;
            for_each_logit(row, V, vec, [&](float l, int) {
                const float e = expf((l - lmax) * inv_t);
                s += e > x ? e : 0.f;
            });
            s = block_sum(s, sh_f);
            if (s < target) hi = mid; else lo = mid;
        }
        thr = __uint_as_float(hi);
    }

    const unsigned long long seed = seeds ? seeds[b] : 0ull;
    const uint32_t step = steps ? static_cast<uint32_t>(steps[b]) : 0u;
    const uint32_t key = lowbias32(static_cast<uint32_t>(seed) + 0x9E3779B9u * step) ^
                         lowbias32(static_cast<uint32_t>(seed >> 32) ^ 0x85EBCA6Bu);
    ArgMax best = {-INFINITY, 0x7fffffff};
    for_each_logit(row, V, vec, [&](float l, int i) {
        if (thr > 0.f && !(expf((l - lmax) * inv_t) >= thr)) return;
        best = better(best, ArgMax{l * inv_t + gumbel(key, static_cast<uint32_t>(i + idx_offset)), i});
    });
    best = block_argmax(best, sh_am);
    if (threadIdx.x == 0) {
        if (pairs) pairs[b] = make_float2(best.v, __int_as_float((best.i < V ? best.i : am.i) + idx_offset));
        else out_ids[b] = best.i < V ? best.i : am.i;
    }
}

// out_ids[b] = index of the best pair among the `world` shards of row b (highest score, lowest index on ties).
__global__ void sample_merge_kernel(const float2* __restrict__ pairs, int world, int B, int* __restrict__ out_ids) {
    pdl_launch_dependents();
    pdl_wait();
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    ArgMax best = {-INFINITY, 0x7fffffff};
    for (int r = 0; r < world; ++r) {
        const float2 p = pairs[static_cast<long long>(r) * B + b];
        best = better(best, ArgMax{p.x, __float_as_int(p.y)});
    }
    out_ids[b] = best.i;
}

}  // namespace atx

using namespace atx;

extern "C" {

int atx_sample(const float* logits, long long ldl, const float* temperature, const float* top_p,
               const unsigned long long* seeds, const int* steps, int* out_ids, int B, int V, int flags,
               atx_stream_t stream) {
    ATX_REQUIRE(logits && out_ids, "atx_sample: null pointer");
    ATX_REQUIRE(B > 0 && V > 0 && ldl >= V, "atx_sample: bad shape B=%d V=%d ldl=%lld", B, V, ldl);
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = (flags & ATX_FLAG_PDL) ? 1 : 0;
    cfg.gridDim = dim3(static_cast<unsigned>(B));
    cfg.blockDim = dim3(kSampThreads);
    cfg.stream = reinterpret_cast<cudaStream_t>(stream);
    ATX_CUDA(cudaLaunchKernelEx(&cfg, sample_kernel, logits, ldl, temperature, top_p, seeds, steps, out_ids, V, 0,
                                static_cast<float2*>(nullptr)));
    return ATX_OK;
}

int atx_sample_shard(const float* logits, long long ldl, int idx_offset, const float* temperature,
                     const unsigned long long* seeds, const int* steps, void* pairs, int B, int V_local, int flags,
                     atx_stream_t stream) {
    ATX_REQUIRE(logits && pairs, "atx_sample_shard: null pointer");
    ATX_REQUIRE(B > 0 && V_local > 0 && ldl >= V_local && idx_offset >= 0,
                "atx_sample_shard: bad shape B=%d V_local=%d ldl=%lld offset=%d", B, V_local, ldl, idx_offset);
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = (flags & ATX_FLAG_PDL) ? 1 : 0;
    cfg.gridDim = dim3(static_cast<unsigned>(B));
    cfg.blockDim = dim3(kSampThreads);
    cfg.stream = reinterpret_cast<cudaStream_t>(stream);
    ATX_CUDA(cudaLaunchKernelEx(&cfg, sample_kernel, logits, ldl, temperature, static_cast<const float*>(nullptr), seeds,
                                steps, static_cast<int*>(nullptr), V_local, idx_offset, static_cast<float2*>(pairs)));
    return ATX_OK;
}

int atx_sample_merge(const void* pairs, int world, int B, int* out_ids, int flags, atx_stream_t stream) {
    ATX_REQUIRE(pairs && out_ids && world > 0 && B > 0, "atx_sample_merge: bad arguments");
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = (flags & ATX_FLAG_PDL) ? 1 : 0;
    cfg.gridDim = dim3(static_cast<unsigned>((B + 127) / 128));
    cfg.blockDim = dim3(128);
    cfg.stream = reinterpret_cast<cudaStream_t>(stream);
    ATX_CUDA(cudaLaunchKernelEx(&cfg, sample_merge_kernel, static_cast<const float2*>(pairs), world, B, out_ids));
    return ATX_OK;
}

}  // extern "C"
