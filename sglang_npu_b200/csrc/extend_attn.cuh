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
};

// tcgen05 path (extend_attn_umma.cu); returns B200_ERR_UNSUPPORTED for shapes it does not cover
int launch_extend_umma(const ExtendArgs& a, int max_extend_len, int head_dim, int dtype, cudaStream_t stream);

}  // namespace b200
