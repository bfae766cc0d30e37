// LSE merge of two partial attention states (natural-log LSE, as the reference's
// merge_state / merge_state_v2: sgl-kernel/csrc/attention/cascade.cu:20-55,
// merge_attn_states.cu:31-204).  One thread per (token, head, 8-element pack).
#include "common.cuh"

namespace b200 {

template <typename T>
__global__ void __launch_bounds__(256) merge_state_kernel(T* __restrict__ v_out, float* __restrict__ s_out,
                                                          const T* __restrict__ v_a, const float* __restrict__ s_a,
                                                          const T* __restrict__ v_b, const float* __restrict__ s_b,
                                                          int64_t num_rows, int packs_per_row) {
  const int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (idx >= num_rows * packs_per_row) return;
  const int64_t row = idx / packs_per_row;
  const float sa = s_a[row], sb = s_b[row];
  // +inf is treated as -inf, like merge_attn_states.cu:52-53
  const float pa = isinf(sa) ? -INFINITY : sa;
  const float pb = isinf(sb) ? -INFINITY : sb;
  const float m = fmaxf(pa, pb);
  float wa = 0.f, wb = 0.f, lse = -INFINITY;
  if (m != -INFINITY) {
    const float ea = expf(pa - m), eb = expf(pb - m);
    const float inv = 1.f / (ea + eb);
    wa = ea * inv;
    wb = eb * inv;
    lse = m + logf(ea + eb);
  }
  const int4 a = *reinterpret_cast<const int4*>(v_a + idx * 8);
  const int4 b = *reinterpret_cast<const int4*>(v_b + idx * 8);
  const uint32_t* au = reinterpret_cast<const uint32_t*>(&a);
  const uint32_t* bu = reinterpret_cast<const uint32_t*>(&b);
  int4 o;
  uint32_t* ou = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 fa = unpack2<T>(au[i]);
    const float2 fb = unpack2<T>(bu[i]);
    ou[i] = pack2<T>(fa.x * wa + fb.x * wb, fa.y * wa + fb.y * wb);
  }
  *reinterpret_cast<int4*>(v_out + idx * 8) = o;
  if (s_out != nullptr && (idx - row * packs_per_row) == 0) s_out[row] = lse;
}

}  // namespace b200

B200_API int b200_merge_state(void* v_out, float* s_out, const void* v_a, const float* s_a,
                                const void* v_b, const float* s_b, int64_t num_tokens, int num_heads,
                                int head_dim, int dtype, void* stream) {
  using namespace b200;
  B200_REQUIRE(num_tokens >= 0 && num_heads > 0 && head_dim > 0, "merge_state: bad sizes");
  if (num_tokens == 0) return B200_OK;
  B200_REQUIRE(v_out && v_a && s_a && v_b && s_b, "merge_state: null pointer");
  B200_REQUIRE(head_dim % 8 == 0, "merge_state: head_dim %d must be a multiple of 8", head_dim);
  B200_REQUIRE(((uintptr_t)v_out | (uintptr_t)v_a | (uintptr_t)v_b) % 16 == 0, "merge_state: 16-byte alignment");
  const int64_t rows = num_tokens * num_heads;
  const int packs = head_dim / 8;
  const int64_t total = rows * packs;
  const unsigned blocks = (unsigned)ceil_div(total, 256);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == B200_BF16)
    merge_state_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>((__nv_bfloat16*)v_out, s_out, (const __nv_bfloat16*)v_a, s_a,
                                                            (const __nv_bfloat16*)v_b, s_b, rows, packs);
  else if (dtype == B200_F16)
    merge_state_kernel<__half><<<blocks, 256, 0, st>>>((__half*)v_out, s_out, (const __half*)v_a, s_a,
                                                     (const __half*)v_b, s_b, rows, packs);
  else
    return fail(B200_ERR_UNSUPPORTED, "merge_state: dtype %d not supported", dtype);
  B200_LAUNCH_OK();
  return B200_OK;
}
