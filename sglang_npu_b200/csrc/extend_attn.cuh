// Shared argument block of the extend (prefill) attention kernels (extend_attn.cu: legacy mma.sync tiles,
// extend_attn_umma.cu: tcgen05 / TMEM tiles).
#pragma once

#include "common.cuh"

namespace b200 {

constexpr float kLog2eExt = 1.4426950408889634f;

struct ExtendArgs {
  void* o;
  const void* q;
  const void* k_new;
  const void* v_new;
  const void* k_cache;
  const void* v_cache;
  const int32_t* req_to_token;
  int64_t req_to_token_stride;
  const void* req_pool_indices;
  int index_is_int64;
  const int32_t* qo_indptr;
  const int32_t* prefix_lens;
  int batch, num_q_heads, num_kv_heads;
  int64_t q_token_stride, o_token_stride, knew_token_stride, vnew_token_stride, kv_token_stride;
  float scale_log2, sm_scale, logit_cap;
  int causal;
  // Segment mode (shared-prefix decode, decode_attn.cu): part_o != NULL.  blockIdx.z = segment s of request 0's prefix,
  // KV positions [s * seg_len, min(prefix, (s+1) * seg_len)) only (no new tokens, no causal mask); every segment uses
  // the same q rows (request 0's).  Instead of o the kernel writes the normalised fp32 partial and its log2-domain
  // LSE into the split-KV workspace of the decode merge: slot s of [row, head, part_stride].
  float* part_o = nullptr;
  float* part_lse = nullptr;
  int part_stride = 0, seg_len = 0;
  int seg_rows = 0, seg_prefix = 0;  // q rows (= decode batch) and shared-prefix length: passed by value, no device arrays
};

// two-query-tile TMA kernel (extend_attn_fa.cu); B200_ERR_UNSUPPORTED (no error text) for what it does not cover
int launch_extend_fa(const ExtendArgs& a, int max_extend_len, int head_dim, int dtype, int64_t num_q_tokens,
                     int64_t num_kv_slots, cudaStream_t stream);

// tcgen05 path (extend_attn_umma.cu); returns B200_ERR_UNSUPPORTED for shapes it does not cover
int launch_extend_umma(const ExtendArgs& a, int max_extend_len, int head_dim, int dtype, cudaStream_t stream);

}  // namespace b200
