/*
 * b200_hotpath.h — C-ABI of the B200-native (sm_100a) SGLang per-layer hot path.
 *
 * This is the drop-in boundary: every entry point replaces one
 * `torch.ops.sgl_kernel.*` op (or Triton kernel) that the reference's Python layers
 * call on this path.  All pointers are raw DEVICE pointers owned by the caller
 * (PyTorch's caching allocator in practice); sizes/strides are in ELEMENTS unless
 * a name says bytes; `stream` is a `cudaStream_t` passed as `void*` (0 = legacy
 * default stream).  Calls only enqueue work on `stream` — no host sync, no
 * allocation — so all of them are CUDA-graph capturable.
 *
 * Return value: 0 (B200_OK) on success, otherwise a b200_status code;
 * `b200_last_error()` returns a thread-local human-readable message.  The Python
 * host layer (sglang_npu_b200/_cabi.py) turns non-zero codes into RuntimeError /
 * NotImplementedError, mirroring TORCH_CHECK / TORCH_CHECK_NOT_IMPLEMENTED in the
 * reference (sgl-kernel/csrc/gemm/int8_gemm_kernel.cu:670-695, :742-743).
 *
 * Reference citations are paths under /root/reference.
 */
#ifndef B200_HOTPATH_H_
#define B200_HOTPATH_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  B200_OK = 0,
  B200_ERR_INVALID_ARG = 1, /* shape / alignment / null-pointer violation      -> RuntimeError        */
  B200_ERR_UNSUPPORTED = 2, /* valid request this build has no kernel for      -> NotImplementedError */
  B200_ERR_CUDA = 3,        /* CUDA runtime / driver error at launch           -> RuntimeError        */
  B200_ERR_WORKSPACE = 4    /* caller-provided workspace too small             -> RuntimeError        */
} b200_status;

typedef enum {
  B200_BF16 = 0,
  B200_F16 = 1,
  B200_F32 = 2,
  B200_FP8_E4M3 = 3,
  B200_I8 = 4
} b200_dtype;

/* ---- library info ------------------------------------------------------------------- */
int b200_abi_version(void);
const char* b200_last_error(void);
/* Fills SM count and compute capability of the current device. */
int b200_device_info(int* sm_count, int* cc_major, int* cc_minor);
/* Diagnostics: when a device buffer is set, the GEMM kernels record per-CTA phase timestamps into it
 * (16 x uint64 per CTA: [0]/[15] = %globaltimer ns at entry/exit, [1..14] = %clock64 at kernel-specific
 * points; CTAs beyond bytes/128 record nothing).  NULL (the default) disables it.  Not thread-safe. */
int b200_debug_set_trace(void* device_buffer, size_t bytes);

/* ---- KV-cache write -------------------------------------------------------------------
 * Replaces MHATokenToKVPool.set_kv_buffer's two index_put_ launches
 * (python/sglang/srt/mem_cache/memory_pool.py:369-407):
 *     k_cache[loc[t], :, :] = k[t, :, :];  v_cache[loc[t], :, :] = v[t, :, :]
 * Bit-exact copy.  k/v rows may be strided views (token stride given, [heads, dim]
 * contiguous); cache rows are [num_kv_heads, head_dim] contiguous with the given
 * token stride.  elem_bytes is the KV element size (2 = bf16/fp16, 1 = fp8 stored
 * as uint8, memory_pool.py:114-118).  Row bytes must be a multiple of 16.
 */
int b200_set_kv_buffer(void* k_cache, void* v_cache, const void* k, const void* v,
                       const void* loc, int loc_is_int64, int64_t num_tokens,
                       int64_t k_row_elems, int64_t v_row_elems,
                       int64_t k_token_stride, int64_t v_token_stride,
                       int64_t k_cache_token_stride, int64_t v_cache_token_stride,
                       int elem_bytes, void* stream);

/* FP8 (e4m3) KV cache, --kv-cache-dtype fp8_e4m3 (memory_pool.py:114-118: stored as uint8): set_kv_buffer's cast branch
 * (memory_pool.py:385-391) fused with the scatter.  k/v are bf16 or f16 (src_dtype); when k_scale / v_scale > 0 the
 * row is first divided by it in fp32 and rounded back to the source dtype (Tensor.div_ by a Python float on the oracle's
 * CPU path), then cast like c10::Float8_e4m3fn(float): round-to-nearest-even, |x| > 464 or NaN -> NaN code (not
 * saturating).  Bit-exact.  The caller's k/v are NOT modified (the reference divides them in place).  Strides in
 * elements; cache rows are [num_kv_heads, head_dim] bytes.  Row elements must be a multiple of 16. */
int b200_set_kv_buffer_fp8(void* k_cache, void* v_cache, const void* k, const void* v, const void* loc,
                           int loc_is_int64, int64_t num_tokens, int64_t k_row_elems, int64_t v_row_elems,
                           int64_t k_token_stride, int64_t v_token_stride, int64_t k_cache_token_stride,
                           int64_t v_cache_token_stride, float k_scale, float v_scale, int src_dtype, void* stream);

/* FP8 KV cache, prefill side: up-cast the cached prefix of every request of an extend batch into a compact 16-bit
 * staging pool: k_out[offsets[b] + pos, :] = T(e4m3(k_cache[req_to_token[req_pool_indices[b], pos], :]) * k_scale) for
 * pos < prefix_lens[b] (same for v).  b200_extend_attn then reads the prefix from the staging pool (page table
 * row b = offsets[b] + 0..prefix) and the new tokens from k_new / v_new — what the reference's Triton extend kernel
 * does with an fp8 k_buffer and bf16 k_extend (triton_ops/extend_attention.py:302-440).  Exact for scale 1. */
int b200_fp8_kv_gather_dequant(void* k_out, void* v_out, const void* k_cache, const void* v_cache,
                               const int32_t* req_to_token, int64_t req_to_token_stride, const void* req_pool_indices,
                               int index_is_int64, const int32_t* prefix_lens, const int32_t* offsets, int batch,
                               int max_prefix_len, int64_t row_elems, int64_t cache_token_stride, float k_scale,
                               float v_scale, int out_dtype, void* stream);

/* ---- decode attention -----------------------------------------------------------------
 * Replaces decode_attention_fwd (python/sglang/srt/layers/attention/triton_ops/
 * decode_attention.py:677-728; stage1 :241-401, stage2 :492-552) and the oracle's
 * _run_sdpa_forward_decode (torch_native_backend.py:112-180):
 *   o[b,h,:] = softmax(sm_scale * q[b,h,:] . K[idx_b, h/g, :]^T) . V[idx_b, h/g, :]
 *   idx_b    = req_to_token[req_pool_indices[b], 0:seq_lens[b]]      (no causal mask)
 * Split-KV: request b is cut into ceil(seq_len_b / kv_chunk) chunks; grid.x covers
 * max_splits chunks, surplus CTAs exit.  Partials (fp32 o, lse) live in `workspace`
 * and are merged by a second kernel.  A request with seq_len 0 produces zeros.
 */
int b200_decode_attn_workspace_bytes(int batch, int num_q_heads, int head_dim, int max_splits,
                                     size_t* bytes);
int b200_decode_attn(void* o, const void* q, const void* k_cache, const void* v_cache,
                     const int32_t* req_to_token, int64_t req_to_token_stride,
                     const void* req_pool_indices, const void* seq_lens, int index_is_int64,
                     int batch, int num_q_heads, int num_kv_heads, int head_dim,
                     int64_t q_token_stride, int64_t o_token_stride,
                     int64_t kv_token_stride, float sm_scale, float logit_cap,
                     int kv_chunk, int max_splits, void* workspace, size_t workspace_bytes,
                     int dtype, void* stream);
/* Same, with stage 2 fused with the per-token int8 quantisation feeding o_proj:
 * (oq, oq_scale) = per_token_quant_int8(o) — bit-identical to b200_decode_attn followed by
 * b200_per_token_quant_int8 on the [batch, num_q_heads * head_dim] output; o is still written. */
int b200_decode_attn_int8(void* o, int8_t* oq, float* oq_scale, const void* q, const void* k_cache,
                          const void* v_cache, const int32_t* req_to_token, int64_t req_to_token_stride,
                          const void* req_pool_indices, const void* seq_lens, int index_is_int64, int batch,
                          int num_q_heads, int num_kv_heads, int head_dim, int64_t q_token_stride,
                          int64_t o_token_stride, int64_t kv_token_stride, float sm_scale, float logit_cap,
                          int kv_chunk, int max_splits, void* workspace, size_t workspace_bytes, int dtype,
                          void* stream);

/* Decode attention over an FP8 (e4m3) KV cache.  K and V are taken at their exact e4m3 values, up-cast
 * before the dot like the reference's Triton kernel does (decode_attention.py:336 `k.to(q.dtype)`); k_scale / v_scale
 * (1.0 = the reference Triton path, which ignores them on the read side) multiply the logits and the output, undoing
 * set_kv_buffer's division.  q / o are `dtype` (bf16 / f16); kv_token_stride in bytes (= elements).  oq / oq_scale
 * may be NULL (plain output) or request the fused o_proj int8 quantisation of b200_decode_attn_int8.  Halves the
 * algorithmic bytes of the step: sum_b len_b * 2 * Hkv * D * 1. */
int b200_decode_attn_fp8kv(void* o, int8_t* oq, float* oq_scale, const void* q, const void* k_cache,
                           const void* v_cache, const int32_t* req_to_token, int64_t req_to_token_stride,
                           const void* req_pool_indices, const void* seq_lens, int index_is_int64, int batch,
                           int num_q_heads, int num_kv_heads, int head_dim, int64_t q_token_stride,
                           int64_t o_token_stride, int64_t kv_token_stride, float sm_scale, float logit_cap,
                           float k_scale, float v_scale, int kv_chunk, int max_splits, void* workspace,
                           size_t workspace_bytes, int dtype, void* stream);

/* Shared-prefix (cascade) decode (the reference's building blocks are merge_state,
 * sgl-kernel/csrc/attention/cascade.cu:20-55, and FlashInfer's cascade wrappers).  Every request of the batch starts
 * with the same `shared_prefix_len` physical KV slots (a radix-cache hit on a common system prompt: BASELINE config 5);
 * request 0's page-table row names them.  Result = b200_decode_attn's (same formula), computed as
 *   1. prefix pass: all batch x Hq query rows against the shared tokens as ONE tensor-core attention problem
 *      (tcgen05, 128 query rows per CTA), cut into segments of `seg_len` tokens (multiple of 128) for SM occupancy;
 *      the shared KV bytes are read once per 128 requests instead of once per request;
 *   2. the split-KV decode kernel over each request's own tokens [shared_prefix_len, seq_len);
 *   3. one LSE merge over the segment partials and the decode partials (the decode stage-2 kernel).
 * Requires seq_lens[b] > shared_prefix_len for every b.  Workspace: b200_decode_attn_workspace_bytes(batch, Hq, D,
 * ceil(shared_prefix_len / seg_len) + max_splits).  oq / oq_scale as in b200_decode_attn_int8, or NULL.
 * num_kv_slots: rows of k_cache / v_cache (TMA extent for the two-query-tile prefix kernel); 0 = unknown. */
int b200_decode_attn_shared_prefix(void* o, int8_t* oq, float* oq_scale, const void* q, const void* k_cache,
                                   const void* v_cache, const int32_t* req_to_token, int64_t req_to_token_stride,
                                   const void* req_pool_indices, const void* seq_lens, int index_is_int64, int batch,
                                   int num_q_heads, int num_kv_heads, int head_dim, int64_t q_token_stride,
                                   int64_t o_token_stride, int64_t kv_token_stride, float sm_scale, float logit_cap,
                                   int shared_prefix_len, int seg_len, int kv_chunk, int max_splits, void* workspace,
                                   size_t workspace_bytes, int dtype, void* stream, int64_t num_kv_slots);

/* Shared-prefix decode over an FP8 (e4m3) KV cache (config 5 with --kv-cache-dtype fp8_e4m3, memory_pool.py:385-391).
 * The caller first up-casts the shared tokens into a 16-bit staging copy with b200_fp8_kv_gather_dequant
 * (k_stage / v_stage [shared_prefix_len, Hkv, D] of `dtype`, values already multiplied by k_scale / v_scale);
 * stage_table[0 .. shared_prefix_len) names its rows (normally 0, 1, 2, ...) and *stage_index (same integer type as
 * req_pool_indices) is 0.  Pass 1 (tensor cores) reads the staging copy, pass 2 (b200_decode_attn_fp8kv's kernel) the
 * e4m3 cache from token shared_prefix_len on; one merge.  Same result as b200_decode_attn_fp8kv. */
int b200_decode_attn_shared_prefix_fp8kv(
    void* o, int8_t* oq, float* oq_scale, const void* q, const void* k_cache, const void* v_cache, const void* k_stage,
    const void* v_stage, const int32_t* stage_table, const void* stage_index, int64_t stage_token_stride,
    const int32_t* req_to_token, int64_t req_to_token_stride, const void* req_pool_indices, const void* seq_lens,
    int index_is_int64, int batch, int num_q_heads, int num_kv_heads, int head_dim, int64_t q_token_stride,
    int64_t o_token_stride, int64_t kv_token_stride, float sm_scale, float logit_cap, float k_scale, float v_scale,
    int shared_prefix_len, int seg_len, int kv_chunk, int max_splits, void* workspace, size_t workspace_bytes, int dtype,
    void* stream);

/* ---- extend (prefill) attention ------------------------------------------------------
 * Replaces extend_attention_fwd (triton_ops/extend_attention.py:302-440) and the
 * oracle's _run_sdpa_forward_extend (torch_native_backend.py:27-110).  Request b has
 * prefix_lens[b] cached tokens and extend_lens[b] = qo_indptr[b+1]-qo_indptr[b] new
 * ones; query i (0-based within the request) attends to all prefix slots plus new
 * tokens 0..i (causal) or all new tokens (causal = 0).  Prefix keys come through the
 * page table; new keys come from k_new/v_new when non-NULL (contiguous activations,
 * token stride given) else also through the page table (they were written first).
 * num_q_tokens / num_kv_slots: rows of q and of k_cache / v_cache (the extents a TMA descriptor needs); 0 = unknown,
 * which keeps the launch on the cp.async gather kernel.
 */
int b200_extend_attn(void* o, const void* q, const void* k_new, const void* v_new,
                     const void* k_cache, const void* v_cache,
                     const int32_t* req_to_token, int64_t req_to_token_stride,
                     const void* req_pool_indices, int index_is_int64,
                     const int32_t* qo_indptr, const int32_t* prefix_lens,
                     int batch, int max_extend_len,
                     int num_q_heads, int num_kv_heads, int head_dim,
                     int64_t q_token_stride, int64_t o_token_stride,
                     int64_t knew_token_stride, int64_t vnew_token_stride,
                     int64_t kv_token_stride, float sm_scale, float logit_cap, int causal,
                     int dtype, void* stream, int64_t num_q_tokens, int64_t num_kv_slots);

/* ---- LSE merge ----------------------------------------------------------------------------
 * Replaces merge_state / merge_state_v2 (sgl-kernel/csrc/attention/cascade.cu:20-55,
 * merge_attn_states.cu:31-204): v_out = (v_a e^{s_a-m} + v_b e^{s_b-m}) / (e^{s_a-m}+e^{s_b-m}),
 * s_out = m + log(...), m = max(s_a, s_b).  v_*: [T, H, D] (dtype), s_*: [T, H] fp32.
 * s_out may be NULL.
 */
int b200_merge_state(void* v_out, float* s_out, const void* v_a, const float* s_a,
                     const void* v_b, const float* s_b, int64_t num_tokens, int num_heads,
                     int head_dim, int dtype, void* stream);

/* ---- W8A8-int8 linear ------------------------------------------------------------------
 * b200_per_token_quant_int8 replaces per_token_quant_int8 (python/sglang/srt/layers/
 * quantization/int8_kernel.py:20-83): per row, absmax = max(|x|, 1e-10);
 * scale = absmax/127; xq = rint(x * (127/absmax)) (round-half-even, the oracle's
 * torch.round rule, test/srt/cpu/utils.py:24-32).
 *
 * b200_int8_scaled_mm replaces int8_scaled_mm (sgl-kernel/csrc/gemm/
 * int8_gemm_kernel.cu:663-746): out[m,n] = cast((sum_k a[m,k]*b[n,k]) * scale_a[m]
 * * scale_b[n] + bias[n]).  `a` is [M,K] row-major int8, `b` is the weight as the
 * reference holds it after process_weights_after_loading — logical [K,N] column-major,
 * i.e. N rows of K contiguous int8 (w8a8_int8.py:332-341).  K % 16 == 0, N % 8 == 0.
 * If acc_out != NULL the raw int32 accumulators [M,N] are written too (parity hook:
 * they must be bit-exact).  Workspace (shared with b200_fp8_scaled_mm): b200_int8_scaled_mm_workspace_bytes
 * bytes, zero-initialised ONCE by the caller, self-cleaning afterwards; it carries the per-CTA partial
 * accumulators and flags of the persistent stream-K decode kernel (M <= 256).  One workspace per stream.
 * workspace == NULL selects the cluster split-K kernel, which needs none.
 */
int b200_per_token_quant_int8(int8_t* xq, float* scale, const void* x, int64_t M, int64_t K,
                              int64_t x_row_stride, int dtype, void* stream);
int b200_int8_scaled_mm_workspace_bytes(int64_t M, int64_t N, int64_t K, size_t* bytes);
int b200_int8_scaled_mm(void* out, const int8_t* a, const int8_t* b, const float* scale_a,
                        const float* scale_b, const void* bias, int64_t M, int64_t N, int64_t K,
                        int out_dtype, int32_t* acc_out, void* workspace, size_t workspace_bytes,
                        void* stream);

/* ---- block-scaled FP8 linear -----------------------------------------------------------
 * b200_per_token_group_quant_fp8 replaces sgl_per_token_group_quant_fp8
 * (sgl-kernel/csrc/gemm/per_token_group_quant_8bit.cu) / per_token_group_quant_fp8
 * (fp8_kernel.py:116-297): per (row, group of `group_size`): s = max(absmax, eps)/448,
 * xq = e4m3(clamp(x / s, +-448)); scales written as [M, K/group] with the given strides
 * (row-major or column-major "TMA-aligned" layout).
 *
 * b200_fp8_blockwise_scaled_mm replaces fp8_blockwise_scaled_mm (sgl-kernel/csrc/gemm/
 * fp8_blockwise_gemm_kernel.cu:345-423): out[m,n] = sum_kb (sum_{k in kb} a[m,k] b[n,k])
 * * sa[m,kb] * sb[n/128,kb]; a [M,K] e4m3 row-major, b N rows of K contiguous e4m3;
 * K % 128 == 0, N % 128 == 0 (fp8_utils.py:190-197).  `workspace`: the self-cleaning stream-K
 * workspace of b200_int8_scaled_mm_workspace_bytes(M, ...) (shared with the other scaled GEMMs of the
 * stream); with a null workspace decode sizes take the cluster split-K kernel instead.
 */
int b200_per_token_group_quant_fp8(void* xq, float* scales, const void* x, int64_t M, int64_t K,
                                   int64_t x_row_stride, int group_size, float eps,
                                   int64_t scale_row_stride, int64_t scale_col_stride,
                                   int dtype, void* stream);
int b200_per_token_quant_fp8(void* xq, float* scale, const void* x, int64_t M, int64_t K,
                             int64_t x_row_stride, int dtype, void* stream);
int b200_fp8_blockwise_scaled_mm(void* out, const void* a, const void* b, const float* scale_a,
                                 int64_t sa_row_stride, int64_t sa_col_stride,
                                 const float* scale_b, int64_t sb_row_stride, int64_t sb_col_stride,
                                 const void* bias, int64_t M, int64_t N, int64_t K, int out_dtype,
                                 void* workspace, size_t workspace_bytes, void* stream);
/* per-token x per-channel FP8 (sgl-kernel/csrc/gemm/fp8_gemm_kernel.cu:1110-1145). */
int b200_fp8_scaled_mm(void* out, const void* a, const void* b, const float* scale_a,
                       const float* scale_b, const void* bias, int64_t M, int64_t N, int64_t K,
                       int out_dtype, void* workspace, size_t workspace_bytes, void* stream);

/* ---- AWQ-int4 linear -------------------------------------------------------------------
 * b200_awq_dequantize replaces awq_dequantize (sgl-kernel/csrc/gemm/awq_kernel.cu:126-221):
 * W[k, 8c+i] = (nib(qweight[k,c], ORDER[i]) - nib(qzeros[k/g,c], ORDER[i])) * scales[k/g, 8c+i],
 * ORDER = [0,4,1,5,2,6,3,7], computed in the scale dtype (sub then mul).
 * b200_awq_gemm fuses that dequantisation into out = x @ W (+bias)
 * (python/sglang/srt/layers/quantization/awq.py:401-418) without materialising W.  It reads the weights in a
 * streaming layout produced ONCE by b200_awq_repack (the process_weights_after_loading hook, awq.py:396-399):
 *   qweight_t [ceil(N/256) * ceil(K/128)][256][16] int32 — per (256-channel tile, 128-k block) a 16 KB unit whose rows
 *             are channels and whose words hold 8 consecutive k of that channel in the AWQ nibble order (16-byte chunks
 *             XOR-swizzled with (row >> 1) & 3, zero padding past N and K): the caller allocates
 *             ceil(N/256) * ceil(K/128) * 16384 bytes;
 *   sz        [K/g, N] uint32 — {scale (low 16 bits), magic + zero point (high 16 bits)}.
 * The dequantised values are bit-identical to b200_awq_dequantize's.  `workspace` (b200_awq_gemm_workspace_bytes,
 * 256-byte aligned, zero-initialised ONCE and then owned by the AWQ GEMMs of one stream: it carries a launch epoch
 * across calls) holds the stream-K state: publish flags and partial accumulators of tiles shared between CTAs.
 */
int b200_awq_dequantize(void* w_out, const int32_t* qweight, const void* scales,
                        const int32_t* qzeros, int64_t K, int64_t N, int group_size,
                        int dtype, void* stream);
int b200_awq_repack(int32_t* qweight_t, uint32_t* sz, const int32_t* qweight, const void* scales,
                    const int32_t* qzeros, int64_t K, int64_t N, int group_size, int dtype, void* stream);
int b200_awq_gemm_workspace_bytes(size_t* bytes);
int b200_awq_gemm(void* out, const void* x, const int32_t* qweight_t, const uint32_t* sz, const void* bias,
                  int64_t M, int64_t N, int64_t K, int group_size, int dtype, void* workspace,
                  size_t workspace_bytes, void* stream);

/* ---- elementwise steps either side of the path -----------------------------
 * fused_add_rmsnorm (sgl-kernel/csrc/elementwise/fused_add_rms_norm_kernel.cu), silu_and_mul
 * (activation.cu), neox rope with a cos/sin cache (rope.cu).  residual may be NULL (plain
 * rmsnorm, out-of-place into `out`).  With residual: residual += x; out = norm(residual)*w.
 */
int b200_rmsnorm(void* out, void* residual, const void* x, const void* weight, int64_t M,
                 int64_t H, float eps, int dtype, void* stream);
int b200_silu_and_mul(void* out, const void* x, int64_t M, int64_t H, int dtype, void* stream);
int b200_rope_neox(void* q, void* k, const float* cos_sin_cache, const int64_t* positions,
                   int64_t num_tokens, int num_q_heads, int num_kv_heads, int head_dim,
                   int rot_dim, int64_t q_token_stride, int64_t k_token_stride, int dtype,
                   void* stream);

/* Greedy sampling (python/sglang/srt/layers/sampler.py greedy path, torch.argmax(logits, -1)): out[r] = index of
 * the largest x[r, :V] (lowest index on ties).  Workspace: b200_argmax_workspace_bytes, zero-initialised once by
 * the caller, self-cleaning; NULL restricts the kernel to one CTA per row. */
int b200_argmax_workspace_bytes(int64_t rows, size_t* bytes);
int b200_argmax(int64_t* out, const void* x, int64_t rows, int64_t V, int64_t row_stride, int dtype,
                void* workspace, size_t workspace_bytes, void* stream);
/* b200_argmax without a caller workspace and without ANY allocation (no cudaMalloc / cudaFree, no allocating library
 * call): the per-row scratch lives in a file-scope `static __device__` array inside the library, 64 KB per GPU, which
 * bounds the call to rows <= 4096 (more rows -> B200_ERR_WORKSPACE; use b200_argmax).  Same result as b200_argmax.
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (CUDA-graph capture, tight serving loops, harnesses that forbid it).  All calls on one GPU share the scratch (it is
 * cleared on `stream` before every launch): calls that overlap on the GPU race — serialise them on one stream. */
int b200_argmax_static(int64_t* out, const void* x, int64_t rows, int64_t V, int64_t row_stride, int dtype, void* stream);
/* Vocab-parallel greedy sampling (replaces the all-gather of the full logits, logits_processor.py:430-500, before
 * sampler.py's argmax): b200_argmax_key stores, per row, a non-negative int64 key = (order-preserving value bits,
 * ~(index_offset + local index)); the TP ranks combine keys with ONE max all-reduce of 8 bytes per row
 * (torch.distributed) and b200_argmax_key_decode turns the winning key into the token id
 * (equal maxima resolve to the lowest global index, like torch.argmax over the gathered logits). */
int b200_argmax_key(int64_t* key_out, const void* x, int64_t rows, int64_t V, int64_t row_stride, int dtype,
                    int64_t index_offset, void* workspace, size_t workspace_bytes, void* stream);
int b200_argmax_key_decode(int64_t* out, const int64_t* keys, int64_t rows, void* stream);
/* hidden[t, :] = table[ids[t], :] (VocabParallelEmbedding.forward, python/sglang/srt/layers/vocab_parallel_embedding.py,
 * replicated table) and, optionally, positions[t] = seq_lens[t] - 1 (forward_batch_info.py:960-962) in the same launch. */
int b200_embedding(void* out, const void* table, const int64_t* ids, int64_t T, int64_t H, int64_t vocab, int dtype,
                   int64_t* positions, const int64_t* seq_lens, void* stream);
/* out[M, N] = a[M, K] . b[N, K]^T (+ bias): un-quantised bf16 / f16 weight-streaming GEMM for decode-size M (<= 256) —
 * the lm_head / logits GEMM (logits_processor.py:470-500 `torch.matmul(hidden, lm_head.weight.T)`), tcgen05 kind::f16
 * through the stream-K kernel; `workspace` as for b200_int8_scaled_mm. */
int b200_dense_gemm_nt(void* out, const void* a, const void* b, const void* bias, int64_t M, int64_t N, int64_t K,
                       int dtype, void* workspace, size_t workspace_bytes, void* stream);

/* ---- fused steps ("fused_add_rmsnorm + emit per-token int8 quant for the following GEMM,
 * RoPE + fused KV-cache write, silu_and_mul + quant for down_proj") ---------------------------------------------
 * Each is bit-identical to the two un-fused entry points run back to back; they remove one launch and one
 * activation round-trip per linear at decode sizes.
 *   b200_rmsnorm_quant_int8:      residual (optional, in place) += x; y = T(rmsnorm(residual) * weight);
 *                                 (xq, scale) = per_token_quant_int8(y); y_out (optional) receives y.
 *   b200_silu_and_mul_quant_int8: y = T(T(silu(x[:, :H])) * x[:, H:]); (xq, scale) = per_token_quant_int8(y).
 *   b200_rope_neox_set_kv:        NeoX RoPE on q, k in place (b200_rope_neox), then
 *                                 k_cache[loc] = k, v_cache[loc] = v (b200_set_kv_buffer); caches may be NULL.
 */
int b200_rmsnorm_quant_int8(int8_t* xq, float* scale, void* y_out, void* residual, const void* x,
                            const void* weight, int64_t M, int64_t H, float eps, int dtype, void* stream);
int b200_silu_and_mul_quant_int8(int8_t* xq, float* scale, const void* x, int64_t M, int64_t H, int dtype,
                                 void* stream);

/* The same two fusions for the block-scaled FP8 linears (fp8.py:444-487 quantises the activation per (token, 128)
 * group before the GEMM: fp8_kernel.py:209-297).  xq [M, H] e4m3, scales [M, H / 128] fp32 row-major; bit-identical
 * to b200_rmsnorm / b200_silu_and_mul followed by b200_per_token_group_quant_fp8(group 128, eps = quant_eps).
 * H % 128 == 0.  residual (may be NULL) is updated in place like fused_add_rmsnorm. */
int b200_rmsnorm_group_quant_fp8(void* xq, float* scales, void* residual, const void* x, const void* weight,
                                 int64_t M, int64_t H, float eps, float quant_eps, int dtype, void* stream);
int b200_silu_and_mul_group_quant_fp8(void* xq, float* scales, const void* x, int64_t M, int64_t H, float quant_eps,
                                      int dtype, void* stream);
int b200_rope_neox_set_kv(void* q, void* k, const void* v, const float* cos_sin_cache, const int64_t* positions,
                          void* k_cache, void* v_cache, const void* loc, int loc_is_int64, int64_t num_tokens,
                          int num_q_heads, int num_kv_heads, int head_dim, int rot_dim, int64_t q_token_stride,
                          int64_t k_token_stride, int64_t v_token_stride, int64_t k_cache_token_stride,
                          int64_t v_cache_token_stride, int dtype, void* stream);

/* ---- caller-side page-table kernels ----------------------------------------------------------
 * Bit-exact integer kernels that replace the Triton kernels feeding the path every step:
 *   b200_alloc_extend   alloc_extend_kernel (python/sglang/srt/mem_cache/allocator.py:276-360): for request b with
 *       prefix p_b and length s_b, token j of the extension gets last_loc_b + 1 + j while the old partial page
 *       lasts, then free_pages[first_new_page_b + k / P] * P + k % P; out_indices are concatenated in request
 *       order; *ret_values = (new_pages << 32) | sum_extend.
 *   b200_alloc_decode   alloc_decode_kernel (allocator.py:364-401): one token per request; *ret_values = new_pages.
 *     Fusions (allocator.py:411 "TODO: fuse last_loc into the kernel"): last_loc == NULL reads it from
 *     req_to_token[req_pool_indices[b], prefix_b - 1] (get_last_loc); write_req_to_token != 0 also stores the new
 *     slots into req_to_token[req, prefix_b : s_b] (write_req_to_token_pool_triton).
 *   b200_write_req_to_token_pool  write_req_to_token_pool_triton (managers/schedule_batch.py:1920-1953)
 *   b200_get_last_loc             get_last_loc_kernel / get_last_loc_torch (schedule_batch.py:1972-2027)
 *   b200_compute_position         compute_position_kernel (model_executor/forward_batch_info.py:886-957)
 *   b200_create_kv_indices        create_flashinfer_kv_indices_triton (layers/attention/utils.py:11-47)
 * `*_is_int64` flags give the element type (int32 / int64) of the index arrays next to them; free_pages,
 * out_indices, ret_values and positions are int64, req_to_token / kv_indptr / kv_indices / extend_start_loc int32.
 */
int b200_alloc_extend(int64_t* out_indices, int64_t* ret_values, const void* prefix_lens, const void* seq_lens,
                      const void* last_loc, int lens_is_int64, const int64_t* free_pages, int batch, int page_size,
                      int32_t* req_to_token, int64_t req_to_token_stride, const void* req_pool_indices,
                      int pool_is_int64, int write_req_to_token, void* stream);
int b200_alloc_decode(int64_t* out_indices, int64_t* ret_values, const void* seq_lens, const void* last_loc,
                      int lens_is_int64, const int64_t* free_pages, int batch, int page_size,
                      int32_t* req_to_token, int64_t req_to_token_stride, const void* req_pool_indices,
                      int pool_is_int64, int write_req_to_token, void* stream);
int b200_write_req_to_token_pool(int32_t* req_to_token, int64_t req_to_token_stride, const void* req_pool_indices,
                                 int pool_is_int64, const void* prefix_lens, const void* seq_lens,
                                 const void* extend_lens, int lens_is_int64, const void* out_cache_loc,
                                 int loc_is_int64, int batch, void* stream);
int b200_get_last_loc(void* result, const int32_t* req_to_token, int64_t req_to_token_stride,
                      const void* req_pool_indices, int pool_is_int64, const void* prefix_lens, int lens_is_int64,
                      int n, void* stream);
int b200_compute_position(int64_t* positions, int32_t* extend_start_loc, const void* extend_prefix_lens,
                          const void* extend_seq_lens, int lens_is_int64, int batch, void* stream);
int b200_create_kv_indices(int32_t* kv_indices, const int32_t* req_to_token, int64_t req_to_token_stride,
                           const void* req_pool_indices, int pool_is_int64, const void* page_kernel_lens,
                           int lens_is_int64, const int32_t* kv_indptr, const int32_t* kv_start_idx, int batch,
                           void* stream);

/* ---- tensor-parallel exchange fused with the following norm + quant --------------------------
 * Replaces, for the row-parallel linears (o_proj, down_proj): tensor_model_parallel_all_reduce
 * (python/sglang/srt/layers/linear.py:1303-1304, distributed/communication_op.py:11-13; custom one-shot kernel
 * sgl-kernel/csrc/allreduce/custom_all_reduce.cuh) + fused_add_rmsnorm + per_token_quant_int8 of the next linear
 * (the reference's partial fusion: layers/flashinfer_comm_fusion.py).  Symmetric buffers are plain device
 * allocations shared with cudaIpc: b200_comm_malloc (zero-filled) / b200_comm_ipc_handle on the owner,
 * b200_comm_ipc_open on the peers.
 *
 * b200_tp_reduce_rmsnorm_quant_int8: `partial_ptrs[r]` = rank r's [M, H] partial GEMM output (as mapped in THIS
 * process; r == tp_rank is the local one).  This rank reduces rows [tp_rank * ceil(M/tp), ...) over all ranks in
 * rank order (fp32, one rounding), adds them to `residual` rows (local, optional, updated in place), applies
 * RMSNorm(weight, eps) and writes, for those rows, into EVERY rank's buffers: int8 codes + scales
 * (xq_ptrs / xscale_ptrs, optional) and/or the normalised activation (y_ptrs, optional).  It returns (stream-
 * ordered) when every rank's rows have landed locally.  flag_ptrs[r]: rank r's 2 * 8 * max_blocks uint32 flag
 * words; epoch: local max_blocks uint32; both zero-initialised once.  All ranks must issue the same sequence
 * of calls.  Arithmetic = b200_rmsnorm_quant_int8 applied to the all-reduced tensor, bit for bit.
 * err_flag: one local zero-initialised uint32; a peer that does not reach a barrier within B200_TP_TIMEOUT_MS
 * (default 20000) makes the kernel record the failure there and finish (no __trap): see b200_comm_check.
 */
int b200_comm_malloc(void** dev_ptr, size_t bytes);
int b200_comm_free(void* dev_ptr);
int b200_comm_ipc_handle(void* dev_ptr, unsigned char* handle64);
int b200_comm_ipc_open(const unsigned char* handle64, void** peer_ptr);
int b200_comm_ipc_close(void* peer_ptr);
int b200_tp_reduce_rmsnorm_quant_int8(const void* const* partial_ptrs, void* const* xq_ptrs,
                                      void* const* xscale_ptrs, void* const* y_ptrs, void* const* flag_ptrs,
                                      void* epoch, void* residual, const void* weight, int64_t M, int64_t H,
                                      float eps, int tp_size, int tp_rank, int max_blocks, int dtype,
                                      void* err_flag, void* stream);
/* Plain all-reduce(sum) of the ranks' [M, H] partials over the same symmetric buffers and flag protocol: the exchange
 * above without residual, norm or quantisation (two-shot: this rank sums its token slice over all ranks in rank order
 * — fp32, one rounding, identical bits on every rank — and stores it into every rank's `out_ptrs[r]`).  Replaces
 * tensor_model_parallel_all_reduce where no fusion applies (python/sglang/srt/distributed/communication_op.py:11-13 ->
 * parallel_state.py:463-542 GroupCoordinator.all_reduce; its small-message kernel is the one/two-shot custom all-reduce,
 * sgl-kernel/csrc/allreduce/custom_all_reduce.cuh:417-453; binding analogue pynccl.py:126-148).  `out_ptrs` may not
 * alias `partial_ptrs`.  Same flag words / epoch / err_flag and the same call-sequence rule as above. */
int b200_tp_all_reduce(const void* const* partial_ptrs, void* const* out_ptrs, void* const* flag_ptrs, void* epoch,
                       int64_t M, int64_t H, int tp_size, int tp_rank, int max_blocks, int dtype, void* err_flag,
                       void* stream);
/* Test entry: all `tp_size` ranks of ONE exchange in a single cooperative launch on one device (blockIdx.y = rank; the
 * pointer tables hold local addresses; epoch / residual / err are per rank).  `args_scratch`: >= tp_size * 512 bytes of
 * device memory.  Lets a single-GPU box execute the exchange kernel's protocol and arithmetic. */
int b200_tp_exchange_loopback(const void* const* partial_ptrs, void* const* xq_ptrs, void* const* xscale_ptrs,
                              void* const* y_ptrs, void* const* flag_ptrs, void* const* epoch_ptrs,
                              void* const* residual_ptrs, const void* weight, int64_t M, int64_t H, float eps,
                              int tp_size, int max_blocks, int dtype, void* const* err_ptrs, void* args_scratch,
                              size_t args_scratch_bytes, void* stream);
/* Synchronises `stream`, then reports (B200_ERR_CUDA + message) a barrier timeout recorded in `err_flag` by an
 * exchange kernel of this rank; B200_OK when every exchange so far met its peers.  The reference's analogue is the
 * NCCL watchdog error surfaced on the next synchronising call. */
int b200_comm_check(const void* err_flag, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_HOTPATH_H_ */
