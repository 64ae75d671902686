/* atx_b200 — C ABI of the B200-native Llama inference kernels (libatx_b200.so).
 *
 * The reference (dlamagna/agentic-traffic-testing) has no FFI of its own: its L5 llm-backend calls
 * the third-party engines vLLM (llm/serve_llm.py:374-379, 521-531) and HF transformers
 * (llm/hf_cpu_server.py:39-50, 91). This header is the seam a replacement engine binds beneath
 * `AsyncVLLMBackend.generate` (llm/serve_llm.py:504): each entry point names the reference
 * arithmetic it replaces. All pointers are DEVICE pointers unless stated; every buffer is owned by
 * the caller; calls are asynchronous on `stream`, never synchronise the host and (unless stated)
 * allocate nothing. Return value: 0 on success, negative on error with text in atx_last_error().
 *
 * Numeric conventions: activations entering a GEMM are bf16, the residual stream and GEMM
 * accumulators are fp32, the KV cache is bf16, softmax / norms reduce in fp32.
 * Index conventions: token ids, positions, slots, block tables and lengths are int32.
 */
#ifndef ATX_B200_H
#define ATX_B200_H

#ifdef __cplusplus
extern "C" {
#endif

typedef void* atx_stream_t; /* a cudaStream_t */

#define ATX_VERSION 100

/* error codes */
#define ATX_OK 0
#define ATX_ERR_ARG -1
#define ATX_ERR_CUDA -2
#define ATX_ERR_UNSUPPORTED -3
#define ATX_ERR_LIMIT -4

/* flags */
#define ATX_FLAG_PDL 1 /* launch with programmatic dependent launch (stream serialization) */
/* atx_gemm_run, ATX_GEMM_ADD_F32, T <= 64: sum the K-splits in a FIXED order (the splits of a tile form a
 * thread-block cluster reduced through distributed shared memory, one CTA adds each output element) instead
 * of letting every split race its red.add: results are bit-reproducible run to run; splits become 1/2/4/8. */
#define ATX_FLAG_ORDERED 2
/* atx_rmsnorm only: `add` points to bf16 [T, H] instead of fp32 (a partial projection all-reduced in bf16) */
#define ATX_FLAG_ADD_BF16 4

/* Thread-local text of the last error raised on the calling thread. Never NULL. */
const char* atx_last_error(void);
int atx_version(void);
/* Device facts of the current device (host pointers may be NULL). */
int atx_device_info(int* sm_count, int* cc_major, int* cc_minor);

/* ---- GEMM: out[t, n] (+)= sum_k X[t, k] * W[n, k] -------------------------------------------------
 * Replaces torch.nn.functional.linear in transformers/models/llama/modeling_llama.py:182-184
 * (gate/up/down), :259-289 (q/k/v/o) and :499 (lm_head). W: [N, K] bf16 row-major (HF layout),
 * X: [T_cap, ldx] bf16. A plan caches the TMA descriptors for one (W, X buffer) pair; creating it
 * is host-only work (no device allocation, no launch). T may vary per call (1..T_cap). */
typedef struct atx_gemm_plan atx_gemm_plan;
#define ATX_GEMM_STORE_F32 0  /* out fp32 = acc                                            */
#define ATX_GEMM_ADD_F32 1    /* out fp32 += acc (atomic; enables split-K; fuses residual) */
#define ATX_GEMM_STORE_BF16 2 /* out bf16 = round(acc)                                     */
/* Fused SwiGLU (modeling_llama.py:182-184): W holds gate and up rows INTERLEAVED in groups of 64
 * (rows [128j, 128j+64) = gate[64j..], rows [128j+64, 128j+128) = up[64j..]); out is bf16
 * [T, N/2] with out[t, 64j+i] = silu(acc_gate) * acc_up. Split-K (T <= 64, splits in {2,4,8} dividing
 * K/64) runs the splits of a tile as one thread-block cluster and sums them through distributed shared
 * memory in a fixed order: no workspace, no atomics, repeatable bit for bit. */
#define ATX_GEMM_SWIGLU_BF16 3
int atx_gemm_plan_create(atx_gemm_plan** plan, const void* W, long long N, long long K, const void* X,
                         long long T_cap, long long ldx);
void atx_gemm_plan_destroy(atx_gemm_plan* plan);
/* splits: K partitions (>=1); 0 = choose (ADD and SWIGLU modes may split). ldo: out row pitch in elements. */
int atx_gemm_run(const atx_gemm_plan* plan, void* out, long long ldo, int T, int mode, int splits,
                 int flags, atx_stream_t stream);
int atx_gemm_auto_splits(long long N, long long K, int T);

/* ---- embedding gather: out[t, :] = float(table[ids[t], :]) ---------------------------------------
 * Replaces LlamaModel.embed_tokens (modeling_llama.py:375-427). table: [V, H] bf16, out fp32. */
int atx_embed_gather(const int* ids, const void* table, float* out, int T, int H, int V,
                     atx_stream_t stream);

/* ---- RMSNorm: out[t, :] = bf16( w * x[r, :] * rsqrt(mean(x[r, :]^2) + eps) ), r = rows ? rows[t] : t
 * Replaces LlamaRMSNorm.forward (modeling_llama.py:62-67) and, with `add`, the residual add before it
 * (LlamaDecoderLayer.forward :313-332). x fp32 (the residual stream), w bf16, out bf16.
 * add (nullable, fp32 [T, H]): x[r, :] += add[r, :] first; without `rows` the sum is written back to x
 * and add is cleared (tensor parallelism: add = all-reduced partial o_proj / down_proj output). With
 * ATX_FLAG_ADD_BF16, add is bf16 [T, H] and is left as it is.
 * Optionally zero-fills `zero_ptr[0 .. zero_bytes)` (16-byte multiple) for a following split-K GEMM. */
int atx_rmsnorm(float* x, float* add, const int* rows, const void* w, void* out, int T, int H, float eps,
                void* zero_ptr, long long zero_bytes, int flags, atx_stream_t stream);

/* ---- tensor-parallel: all-reduce + residual add + RMSNorm in one kernel over NVLink peer memory -----
 * Replaces the NCCL all-reduce vLLM issues after o_proj / down_proj of a tensor-parallel Llama layer
 * (the reference reaches it through llm/serve_llm.py:374-379) together with the residual add and
 * LlamaRMSNorm that follow (modeling_llama.py:313-332, 62-67). One process per GPU; every rank calls
 * it with the same T. Two forms, both two-shot with rank r owning the token rows t = r (mod world).
 *
 * Flag-barrier form (up to 256 rows).  part_ptrs[p] / flag_ptrs[p] (HOST arrays of `world` device pointers) are
 * rank p's partial-sum buffer [T, H] fp32 and flag array int[2 * world], both mapped into this process (symmetric /
 * IPC memory).  x fp32 is the residual stream, w the bf16 norm weight, `rows` (nullable) gathers rows before
 * the norm.  clear_buf (nullable): the OTHER ping-pong partial buffer of this rank, first clear_rows rows are
 * zeroed after the cross-GPU barrier.  state: int[2] in local device memory, zero-initialised once.
 * zero_ptr / zero_bytes as in atx_rmsnorm.  The rank pulls
 * only its own rows of every peer's partial buffer, keeps the fp32 residual of its own rows in x (rows it does
 * not own go stale: x must not be read whole between calls), and writes the normalised bf16 rows into EVERY
 * rank's activation buffer through out_ptrs[p] (HOST array of `world` peer-mapped device pointers to the same
 * [T, H] bf16 buffer on each rank). A second cross-GPU barrier inside the kernel keeps it running until
 * every peer's rows have arrived here, so the consumer only needs stream order. flag_ptrs[p] must hold
 * int[2 * world]. NVLink bytes per rank: (world-1)/world * T*H*6. */
int atx_reduce_scatter_rmsnorm_all_gather(float* x, const void* const* part_ptrs, void* const* flag_ptrs,
                                          void* const* out_ptrs, void* clear_buf, int clear_rows, int* state, int rank,
                                          int world, const int* rows, const void* w, int T, int H, float eps,
                                          void* zero_ptr, long long zero_bytes, int flags, atx_stream_t stream);

/* Low-latency variant for decode-sized steps: no flag barrier and no pulls. Every buffer a peer writes into is
 * kept filled with a sentinel bit pattern (-0.0: fp32 0x80000000, bf16 0x8000; payloads are sanitised so they
 * never contain it) and the receiver spins on the data itself. Ownership of rows as above (t mod world).
 *   part        this rank's partial sums [T, H] fp32 (LOCAL memory; read only)
 *   rs_ptrs[p]  rank p's receive slots for THIS parity, fp32 [world][T_cap][H], peer-mapped, sentinel-filled once
 *               at start-up (the kernel re-arms what it reads)
 *   out_ptrs[p] rank p's activation buffer for THIS parity, bf16 [>= T, H], peer-mapped, sentinel in rows [0, T)
 *               when the call starts
 *   reset_ptr / reset_bytes  LOCAL region filled with the bf16 sentinel: the OTHER parity's activation rows,
 *               which the GEMMs before this call have finished reading (16-byte multiple)
 *   clear_buf / clear_rows, zero_ptr / zero_bytes, rows, x, w  as in the flag-barrier form (local housekeeping)
 * Calls must alternate between two parities (two rs / out buffer sets), every rank making the same calls in
 * the same order. The whole grid (T x ceil(H/2048) CTAs) spins on remote data and must be co-resident: rejected
 * otherwise. Latency per call, 1 MiB message: see profiles/ (ar_microbench). */
int atx_allreduce_add_rmsnorm_ll(float* x, const float* part, void* const* rs_ptrs, void* const* out_ptrs,
                                 void* clear_buf, int clear_rows, void* reset_ptr, long long reset_bytes, int T_cap,
                                 int rank, int world, const int* rows, const void* w, int T, int H, float eps,
                                 void* zero_ptr, long long zero_bytes, int flags, atx_stream_t stream);

/* ---- RoPE + KV-cache write ------------------------------------------------------------------------
 * Replaces apply_rotary_pos_emb (modeling_llama.py:146-168, rotate_half layout :138-143) and
 * past_key_values.update (:267) / vLLM reshape_and_cache. qkv: [T, (Hq + 2*Hkv) * D] fp32 (GEMM
 * output), cos_sin: [max_pos, D] fp32 (row p = cos(p*inv_freq[0..D/2)) ++ sin(...)), positions[T],
 * slots[T] (slot = block * block_size + offset; negative = skip the cache write).
 * q_out: [T, Hq * D] bf16. K and V caches: [num_blocks, Hkv, block_size, D] bf16 each. */
int atx_rope_kv_write(const float* qkv, const int* positions, const int* slots, const float* cos_sin,
                      void* q_out, void* k_cache, void* v_cache, int T, int Hq, int Hkv, int D,
                      int block_size, int flags, atx_stream_t stream);

/* ---- SwiGLU: act[t, i] = bf16( silu(gu[t, i]) * gu[t, I + i] ) ------------------------------------
 * Replaces LlamaMLP.forward's act_fn(gate) * up (modeling_llama.py:182-184). gu: [T, 2*I] fp32. */
int atx_silu_mul(const float* gu, void* act, int T, int I, int flags, atx_stream_t stream);

/* ---- paged attention, decode (one query token per sequence) ---------------------------------------
 * Replaces eager_attention_forward (modeling_llama.py:199-222; repeat_kv :187-196; fp32 softmax
 * :217) over vLLM's paged KV (FlashInfer decode). q: [B, Hq * D] bf16; out: [B, Hq * D] bf16;
 * block_table: [B, max_blocks] int32; seq_lens[B] = tokens in cache INCLUDING the current one.
 * workspace: fp32 scratch of atx_paged_attn_decode_workspace_bytes(B, Hq, D, num_splits) bytes,
 * supplied by the caller (partial outputs + running max / sum per split).
 * head_dim 64 | 128, at most 8 query heads per KV head, block_size a multiple of 16 (else ATX_ERR_UNSUPPORTED). */
long long atx_paged_attn_decode_workspace_bytes(int B, int Hq, int D, int num_splits);
int atx_paged_attn_decode(const void* q, const void* k_cache, const void* v_cache, const int* block_table,
                          const int* seq_lens, void* out, void* workspace, long long workspace_bytes,
                          int B, int Hq, int Hkv, int D, int block_size, int max_blocks, int num_splits,
                          float scale, int flags, atx_stream_t stream);

/* Decode attention for callers that cannot allocate device memory. (No counterpart in the reference:
 * it exists for hosts of this library, not for parity.) Same arithmetic and arguments as
 * atx_paged_attn_decode, but the per-split partial results are kept in an 8 MiB array that is built
 * into the library (resident on every GPU that loads it, until the process exits) instead of in a
 * caller buffer — for processes that have already committed all of their HBM (e.g. to a KV pool),
 * run under a fixed memory reservation, must avoid cudaMalloc's device-wide synchronisation on the
 * request path, or bind this ABI from a language runtime with no CUDA allocator of its own.
 *   - Limit: atx_paged_attn_decode_workspace_bytes(B, Hq, D, num_splits) must not exceed
 *     atx_paged_attn_decode_static_capacity() (8 MiB); larger requests return ATX_ERR_LIMIT with the
 *     sizes spelled out in atx_last_error(). It never calls cudaMalloc / cudaFree.
 *   - The first call on each device looks up the array's device address (one host-side runtime
 *     query, with that device current); make that first call before starting any stream capture.
 *   - All calls on a device share the one array and it keeps whatever the last call left in it.
 *     Nothing needs clearing between calls (each call overwrites every slot it later reads), but two
 *     calls whose kernels overlap on the GPU — different streams, or different host threads — race
 *     on it: serialise them (same stream, or an event) or use the caller-scratch entry point. */
long long atx_paged_attn_decode_static_capacity(void);
int atx_paged_attn_decode_static(const void* q, const void* k_cache, const void* v_cache,
                                 const int* block_table, const int* seq_lens, void* out, int B, int Hq,
                                 int Hkv, int D, int block_size, int max_blocks, int num_splits, float scale,
                                 int flags, atx_stream_t stream);

/* ---- paged attention, prefill / chunked prefill (causal) ------------------------------------------
 * Same arithmetic as above with the causal mask of create_causal_mask; queries of sequence s are
 * rows [cu_q[s], cu_q[s+1]) of q, and their keys are the first ctx_lens[s] cache tokens of that
 * sequence (the chunk's own K/V must already be in the cache): query j of n attends to keys
 * [0, ctx_lens[s] - n + j]. q/out: [total_q, Hq * D] bf16 (total_q = cu_q[S]); num_blocks = blocks in
 * each cache (they bound the TMA descriptors). One kernel, on tcgen05: head_dim 64 | 128, block_size 16,
 * 16-byte aligned buffers; anything else returns ATX_ERR_UNSUPPORTED. */
int atx_paged_attn_prefill(const void* q, const void* k_cache, const void* v_cache, const int* block_table,
                           const int* cu_q, const int* ctx_lens, void* out, int S, int max_q_len,
                           long long total_q, long long num_blocks, int Hq, int Hkv, int D, int block_size,
                           int max_blocks, float scale, int flags, atx_stream_t stream);

/* ---- sampler --------------------------------------------------------------------------------------
 * Replaces vLLM SamplingParams(temperature=0.2) (llm/serve_llm.py:379, 521) / HF do_sample
 * (llm/hf_cpu_server.py:91) and greedy arg-max used for parity. logits: [B, ldl] fp32, V valid.
 * temperature[b] <= 0 => arg-max (lowest index wins ties, as torch.argmax). Otherwise an exact
 * categorical draw from softmax(logits / temperature) restricted to the top-p nucleus, by the
 * Gumbel-max rule with the counter-based generator documented in oracle/sampler_ref.py
 * (seeds[b], steps[b] select the stream). out_ids: [B] int32. */
int atx_sample(const float* logits, long long ldl, const float* temperature, const float* top_p,
               const unsigned long long* seeds, const int* steps, int* out_ids, int B, int V, int flags,
               atx_stream_t stream);

/* Vocabulary-parallel form of the above (tensor parallelism: each rank holds V / world columns of every logits row,
 * the LM head being sharded by vocabulary): atx_sample_shard reduces this rank's shard of each row to ONE
 * (score, global index) pair — the logit for greedy rows, logit / T + Gumbel noise keyed by the GLOBAL token index
 * otherwise — and atx_sample_merge, given the `world` shards' pairs ([world][B] float2: score, index bits), picks the
 * best (lowest index on ties): exactly the token atx_sample draws from the assembled row, with 8 bytes per row and
 * rank exchanged instead of the row itself. Replaces the logits all-gather vLLM performs before its sampler under
 * tensor parallelism (reached through llm/serve_llm.py:374-379). No top-p here (it needs the whole row's sum). */
int atx_sample_shard(const float* logits, long long ldl, int idx_offset, const float* temperature,
                     const unsigned long long* seeds, const int* steps, void* pairs, int B, int V_local, int flags,
                     atx_stream_t stream);
int atx_sample_merge(const void* pairs, int world, int B, int* out_ids, int flags, atx_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* ATX_B200_H */
