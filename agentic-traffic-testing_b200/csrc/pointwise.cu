// Bandwidth-bound glue between the GEMMs: embedding gather, RMSNorm, RoPE + KV-page scatter, SwiGLU.
// All fp32 math; inputs/outputs as documented in include/atx_b200.h. Every kernel starts with
// griddepcontrol.wait so it can be launched with programmatic dependent launch.
#include "common.cuh"

#include "../../include/atx_b200.h"

namespace atx {

// ---- embedding gather (modeling_llama.py:375-427 embed_tokens) -------------------------------------
__global__ void __launch_bounds__(128) embed_gather_kernel(const int* __restrict__ ids,
                                                           const uint4* __restrict__ table,
                                                           float4* __restrict__ out, int H, int V) {
    pdl_launch_dependents();
    pdl_wait();
    const int t = blockIdx.x;
    int id = ids[t];
    id = id < 0 ? 0 : (id >= V ? V - 1 : id);  // clamp: never read outside the table
    const uint4* src = table + static_cast<long long>(id) * (H / 8);
    float4* dst = out + static_cast<long long>(t) * (H / 4);
    for (int i = threadIdx.x; i < H / 8; i += blockDim.x) {
        const uint4 v = __ldg(src + i);
        dst[2 * i] = make_float4(bf16lo(v.x), bf16hi(v.x), bf16lo(v.y), bf16hi(v.y));
        dst[2 * i + 1] = make_float4(bf16lo(v.z), bf16hi(v.z), bf16lo(v.w), bf16hi(v.w));
    }
}

// ---- RMSNorm (modeling_llama.py:62-67) ---------------------------------------------------------------
constexpr int kNormThreads = 256;
constexpr int kNormMaxVec = 8;  // float4 per thread -> H <= 8192

__global__ void __launch_bounds__(kNormThreads)
rmsnorm_kernel(float* __restrict__ x, float* __restrict__ add, const int* __restrict__ rows,
               const uint2* __restrict__ w, uint2* __restrict__ out, int T, int H, float eps,
               uint4* __restrict__ zero_ptr, long long zero_vecs, int add_bf16) {
    __shared__ float red[kNormThreads / 32];
    pdl_launch_dependents();
    pdl_wait();
    if (blockIdx.x >= T) {
        // extra blocks: clear the accumulation buffer of the split-K GEMM that follows
        const long long nb = gridDim.x - T;
        for (long long i = static_cast<long long>(blockIdx.x - T) * blockDim.x + threadIdx.x; i < zero_vecs;
             i += nb * blockDim.x)
            zero_ptr[i] = make_uint4(0, 0, 0, 0);
        return;
    }
    const int t = blockIdx.x;
    const int r = rows ? rows[t] : t;
    float4* xr = reinterpret_cast<float4*>(x + static_cast<long long>(r) * H);
    float4* ar = (add && !add_bf16) ? reinterpret_cast<float4*>(add + static_cast<long long>(r) * H) : nullptr;
    // ATX_FLAG_ADD_BF16: the partial projection was all-reduced in bf16 (prefill-sized tensor-parallel steps, as
    // vLLM does for every step): read only, nothing to hand back cleared
    const uint2* ab = (add && add_bf16) ? reinterpret_cast<const uint2*>(add) + static_cast<long long>(r) * (H / 4) : nullptr;
    const int nv = H / 4;
    float4 v[kNormMaxVec];
    float ss = 0.f;
#pragma unroll
    for (int j = 0; j < kNormMaxVec; ++j) {
        const int i = threadIdx.x + j * kNormThreads;
        if (i < nv) {
            v[j] = xr[i];
            if (ar) {
                // tensor-parallel residual: x += (all-reduced partial projection); the partial buffer is
                // handed back cleared for the next split-K accumulation (not in row-gather mode)
                const float4 a = ar[i];
                v[j].x += a.x; v[j].y += a.y; v[j].z += a.z; v[j].w += a.w;
                if (!rows) {
                    xr[i] = v[j];
                    ar[i] = make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
            if (ab) {
                const uint2 a = ab[i];
                v[j].x += bf16lo(a.x); v[j].y += bf16hi(a.x); v[j].z += bf16lo(a.y); v[j].w += bf16hi(a.y);
                if (!rows) xr[i] = v[j];
            }
            ss += v[j].x * v[j].x + v[j].y * v[j].y + v[j].z * v[j].z + v[j].w * v[j].w;
        }
    }
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    float tot = 0.f;
#pragma unroll
    for (int i = 0; i < kNormThreads / 32; ++i) tot += red[i];
    const float rs = rsqrtf(tot / static_cast<float>(H) + eps);
    uint2* orow = out + static_cast<long long>(t) * nv;
#pragma unroll
    for (int j = 0; j < kNormMaxVec; ++j) {
        const int i = threadIdx.x + j * kNormThreads;
        if (i < nv) {
            const uint2 wv = __ldg(w + i);
            uint2 o;
            o.x = pack_bf16(bf16lo(wv.x) * (v[j].x * rs), bf16hi(wv.x) * (v[j].y * rs));
            o.y = pack_bf16(bf16lo(wv.y) * (v[j].z * rs), bf16hi(wv.y) * (v[j].w * rs));
            orow[i] = o;
        }
    }
}

// ---- RoPE + KV page scatter (modeling_llama.py:138-168, :267) ---------------------------------------
// One block per token. Work items: (q or k head, 4 rotation pairs) and (v head, 4 elements).
__global__ void __launch_bounds__(256)
rope_kv_write_kernel(const float* __restrict__ qkv, const int* __restrict__ positions,
                     const int* __restrict__ slots, const float* __restrict__ cos_sin,
                     __nv_bfloat16* __restrict__ q_out, __nv_bfloat16* __restrict__ k_cache,
                     __nv_bfloat16* __restrict__ v_cache, int Hq, int Hkv, int D, int block_size) {
    pdl_launch_dependents();
    pdl_wait();
    const int t = blockIdx.x;
    const int pos = positions[t];
    const int slot = slots[t];
    const int half = D / 2;
    const int per_head = half / 4;  // work items per rotated head
    const int n_rot = (Hq + Hkv) * per_head;
    const int n_v = Hkv * (D / 4);
    const float* row = qkv + static_cast<long long>(t) * (Hq + 2 * Hkv) * D;
    const float* cs = cos_sin + static_cast<long long>(pos) * D;
    const long long blk = slot >= 0 ? slot / block_size : 0;
    const int off = slot >= 0 ? slot % block_size : 0;
    for (int w = blockIdx.y * blockDim.x + threadIdx.x; w < n_rot + n_v; w += gridDim.y * blockDim.x) {
        if (w < n_rot) {
            const int h = w / per_head;
            const int i = (w % per_head) * 4;
            const float4 x1 = *reinterpret_cast<const float4*>(row + h * D + i);
            const float4 x2 = *reinterpret_cast<const float4*>(row + h * D + half + i);
            const float4 c = *reinterpret_cast<const float4*>(cs + i);
            const float4 s = *reinterpret_cast<const float4*>(cs + half + i);
            // q*cos + rotate_half(q)*sin with separately rounded products, as the reference does
            uint2 lo, hi;
            lo.x = pack_bf16(__fadd_rn(__fmul_rn(x1.x, c.x), __fmul_rn(-x2.x, s.x)),
                             __fadd_rn(__fmul_rn(x1.y, c.y), __fmul_rn(-x2.y, s.y)));
            lo.y = pack_bf16(__fadd_rn(__fmul_rn(x1.z, c.z), __fmul_rn(-x2.z, s.z)),
                             __fadd_rn(__fmul_rn(x1.w, c.w), __fmul_rn(-x2.w, s.w)));
            hi.x = pack_bf16(__fadd_rn(__fmul_rn(x2.x, c.x), __fmul_rn(x1.x, s.x)),
                             __fadd_rn(__fmul_rn(x2.y, c.y), __fmul_rn(x1.y, s.y)));
            hi.y = pack_bf16(__fadd_rn(__fmul_rn(x2.z, c.z), __fmul_rn(x1.z, s.z)),
                             __fadd_rn(__fmul_rn(x2.w, c.w), __fmul_rn(x1.w, s.w)));
            __nv_bfloat16* dst;
            if (h < Hq) {
                dst = q_out + (static_cast<long long>(t) * Hq + h) * D;
            } else {
                if (slot < 0) continue;
                dst = k_cache + ((blk * Hkv + (h - Hq)) * block_size + off) * D;
            }
            *reinterpret_cast<uint2*>(dst + i) = lo;
            *reinterpret_cast<uint2*>(dst + half + i) = hi;
        } else {
            if (slot < 0) continue;
            const int u = w - n_rot;
            const int h = u / (D / 4);
            const int i = (u % (D / 4)) * 4;
            const float4 x = *reinterpret_cast<const float4*>(row + (Hq + Hkv + h) * D + i);
            uint2 o;
            o.x = pack_bf16(x.x, x.y);
            o.y = pack_bf16(x.z, x.w);
            *reinterpret_cast<uint2*>(v_cache + ((blk * Hkv + h) * block_size + off) * D + i) = o;
        }
    }
}

// ---- SwiGLU (modeling_llama.py:182-184) ----------------------------------------------------------------
__global__ void __launch_bounds__(256)
silu_mul_kernel(const float4* __restrict__ gu, uint2* __restrict__ act, long long total_vec, int I4) {
    pdl_launch_dependents();
    pdl_wait();
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total_vec;
         idx += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long t = idx / I4;
        const int i = static_cast<int>(idx - t * I4);
        const float4 g = gu[t * 2 * I4 + i];
        const float4 u = gu[t * 2 * I4 + I4 + i];
        uint2 o;
        o.x = pack_bf16(g.x / (1.f + expf(-g.x)) * u.x, g.y / (1.f + expf(-g.y)) * u.y);
        o.y = pack_bf16(g.z / (1.f + expf(-g.z)) * u.z, g.w / (1.f + expf(-g.w)) * u.w);
        act[idx] = o;
    }
}

static void set_pdl(cudaLaunchConfig_t& cfg, cudaLaunchAttribute* attr, int flags) {
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = (flags & ATX_FLAG_PDL) ? 1 : 0;
}

}  // namespace atx

using namespace atx;

extern "C" {

int atx_embed_gather(const int* ids, const void* table, float* out, int T, int H, int V, atx_stream_t stream) {
    ATX_REQUIRE(ids && table && out, "atx_embed_gather: null pointer");
    ATX_REQUIRE(T > 0 && V > 0 && H > 0 && H % 8 == 0, "atx_embed_gather: bad shape T=%d H=%d V=%d", T, H, V);
    embed_gather_kernel<<<T, 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        ids, reinterpret_cast<const uint4*>(table), reinterpret_cast<float4*>(out), H, V);
    ATX_LAUNCH_CHECK("embed_gather_kernel");
    return ATX_OK;
}

int atx_rmsnorm(float* x, float* add, const int* rows, const void* w, void* out, int T, int H, float eps,
                void* zero_ptr, long long zero_bytes, int flags, atx_stream_t stream) {
    ATX_REQUIRE(x && w && out, "atx_rmsnorm: null pointer");
    ATX_REQUIRE(T > 0 && H > 0 && H % 4 == 0 && H <= kNormThreads * kNormMaxVec * 4,
                "atx_rmsnorm: bad shape T=%d H=%d (H %% 4 == 0, H <= %d)", T, H, kNormThreads * kNormMaxVec * 4);
    ATX_REQUIRE(zero_bytes >= 0 && zero_bytes % 16 == 0 && (zero_bytes == 0 || zero_ptr),
                "atx_rmsnorm: zero_bytes=%lld must be a non-negative multiple of 16 with a buffer", zero_bytes);
    const long long zv = zero_bytes / 16;
    long long zb = (zv + kNormThreads * 8 - 1) / (kNormThreads * 8);
    if (zb > 2048) zb = 2048;
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[1];
    cfg.gridDim = dim3(static_cast<unsigned>(T + zb));
    cfg.blockDim = dim3(kNormThreads);
    cfg.stream = reinterpret_cast<cudaStream_t>(stream);
    set_pdl(cfg, attr, flags);
    ATX_CUDA(cudaLaunchKernelEx(&cfg, rmsnorm_kernel, x, add, rows, reinterpret_cast<const uint2*>(w),
                                reinterpret_cast<uint2*>(out), T, H, eps, reinterpret_cast<uint4*>(zero_ptr), zv,
                                (flags & ATX_FLAG_ADD_BF16) ? 1 : 0));
    return ATX_OK;
}

int atx_rope_kv_write(const float* qkv, const int* positions, const int* slots, const float* cos_sin,
                      void* q_out, void* k_cache, void* v_cache, int T, int Hq, int Hkv, int D,
                      int block_size, int flags, atx_stream_t stream) {
    ATX_REQUIRE(qkv && positions && slots && cos_sin && q_out && k_cache && v_cache,
                "atx_rope_kv_write: null pointer");
    ATX_REQUIRE(T > 0 && Hq > 0 && Hkv > 0 && D > 0 && D % 8 == 0 && block_size > 0,
                "atx_rope_kv_write: bad shape T=%d Hq=%d Hkv=%d D=%d block_size=%d", T, Hq, Hkv, D, block_size);
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[1];
    // few tokens (decode): split each token's (Hq + 2 Hkv) * D / 4 work items over several CTAs
    const int items = (Hq + Hkv) * (D / 8) + Hkv * (D / 4);
    const int per_tok = T >= 512 ? 1 : (items + 255) / 256;
    cfg.gridDim = dim3(static_cast<unsigned>(T), static_cast<unsigned>(per_tok));
    cfg.blockDim = dim3(256);
    cfg.stream = reinterpret_cast<cudaStream_t>(stream);
    set_pdl(cfg, attr, flags);
    ATX_CUDA(cudaLaunchKernelEx(&cfg, rope_kv_write_kernel, qkv, positions, slots, cos_sin,
                                reinterpret_cast<__nv_bfloat16*>(q_out), reinterpret_cast<__nv_bfloat16*>(k_cache),
                                reinterpret_cast<__nv_bfloat16*>(v_cache), Hq, Hkv, D, block_size));
    return ATX_OK;
}

int atx_silu_mul(const float* gu, void* act, int T, int I, int flags, atx_stream_t stream) {
    ATX_REQUIRE(gu && act, "atx_silu_mul: null pointer");
    ATX_REQUIRE(T > 0 && I > 0 && I % 4 == 0, "atx_silu_mul: bad shape T=%d I=%d", T, I);
    const long long total = static_cast<long long>(T) * (I / 4);
    long long blocks = (total + 255) / 256;
    const long long cap = static_cast<long long>(num_sms()) * 16;
    if (blocks > cap) blocks = cap;
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[1];
    cfg.gridDim = dim3(static_cast<unsigned>(blocks));
    cfg.blockDim = dim3(256);
    cfg.stream = reinterpret_cast<cudaStream_t>(stream);
    set_pdl(cfg, attr, flags);
    ATX_CUDA(cudaLaunchKernelEx(&cfg, silu_mul_kernel, reinterpret_cast<const float4*>(gu),
                                reinterpret_cast<uint2*>(act), total, I / 4));
    return ATX_OK;
}

}  // extern "C"
