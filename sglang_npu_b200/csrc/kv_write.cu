// set_kv_buffer: scatter the step's K/V rows into the paged KV pool (bit-exact copy).
// Replaces the two index_put_ launches of MHATokenToKVPool.set_kv_buffer
// (reference python/sglang/srt/mem_cache/memory_pool.py:369-407) with one launch.
//
// HBM-bound byte work: each thread moves one 16-byte chunk of one token row; a warp
// covers 512 contiguous bytes of a row, so both the gather (k/v activations) and the
// scatter (cache rows) are fully coalesced.  blockIdx.y selects K (0) or V (1).
#include "common.cuh"
#include <type_traits>

namespace b200 {

struct KvWriteArgs {
  uint8_t* cache[2];
  const uint8_t* src[2];
  int64_t row_bytes[2];
  int64_t src_stride_bytes[2];
  int64_t cache_stride_bytes[2];
  const void* loc;
  int loc_is_int64;
  int64_t num_tokens;
};

__global__ void __launch_bounds__(256) kv_write_kernel(const KvWriteArgs a) {
  pdl_trigger();
  pdl_wait();
  const bool isv = blockIdx.y != 0;
  uint8_t* const cache = isv ? a.cache[1] : a.cache[0];
  const uint8_t* const src = isv ? a.src[1] : a.src[0];
  const int64_t row_bytes = isv ? a.row_bytes[1] : a.row_bytes[0];
  const int64_t src_stride = isv ? a.src_stride_bytes[1] : a.src_stride_bytes[0];
  const int64_t cache_stride = isv ? a.cache_stride_bytes[1] : a.cache_stride_bytes[0];
  const int64_t chunks_per_row = row_bytes >> 4;
  const int64_t total = a.num_tokens * chunks_per_row;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t t = i / chunks_per_row;
    const int64_t c = i - t * chunks_per_row;
    const int64_t slot = load_index(a.loc, t, a.loc_is_int64);
    const int4 v = ld_nc_v4(src + t * src_stride + (c << 4));
    *reinterpret_cast<int4*>(cache + slot * cache_stride + (c << 4)) = v;
  }
}

// FP8 (e4m3) KV cache: set_kv_buffer's cast branch (memory_pool.py:385-391) fused with the scatter:
//     if scale given: x = T(float(x) / scale)     (Tensor.div_ of a bf16/f16 tensor by a Python float, fp32 math)
//     cache[loc[t]] = float8_e4m3fn(x)            (round-to-nearest-even; |x| > 464 or NaN -> NaN, as c10's cast does,
//                                                  not saturating)
// Each thread converts 16 source elements (two 16-byte loads) into one 16-byte store.
template <typename T>
__device__ __forceinline__ uint32_t to_e4m3_torch(float x, uint32_t sign, float scale, bool has_scale) {
  // `sign` (0x80 or 0) comes from the raw 16-bit pattern: the f16 -> f32 convert drops a NaN's sign
  if (has_scale) {
    // NaN sign after the division, as ATen's CPU kernels leave it (pinned by tests/golden/kvwrite_fp8.pt): the
    // vectorised fp32 -> bf16 pack writes NaN as 0xFFFF (negative); the f16 pack keeps the operand's sign.  The GPU
    // divide would canonicalise it to +NaN.
    if (std::is_same<T, __nv_bfloat16>::value && x != x) sign = 0x80u;
    x = DT<T>::to_f(DT<T>::from_f(__fdiv_rn(x, scale)));
  }
  uint16_t r;
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;\n" : "=h"(r) : "f"(0.f), "f"(x));
  uint32_t b = r & 0xffu;
  if (!(fabsf(x) <= 464.f)) b = 0x7fu | sign;
  return b;
}

template <typename T>
__global__ void __launch_bounds__(256) kv_write_fp8_kernel(const KvWriteArgs a, float k_scale, float v_scale) {
  pdl_trigger();
  pdl_wait();
  const bool isv = blockIdx.y != 0;
  uint8_t* const cache = isv ? a.cache[1] : a.cache[0];
  const uint8_t* const src = isv ? a.src[1] : a.src[0];
  const int64_t row_bytes = isv ? a.row_bytes[1] : a.row_bytes[0];  // of the fp8 cache row
  const int64_t src_stride = isv ? a.src_stride_bytes[1] : a.src_stride_bytes[0];
  const int64_t cache_stride = isv ? a.cache_stride_bytes[1] : a.cache_stride_bytes[0];
  const float scale = isv ? v_scale : k_scale;
  const bool has_scale = scale > 0.f;
  const int64_t chunks_per_row = row_bytes >> 4;
  const int64_t total = a.num_tokens * chunks_per_row;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t t = i / chunks_per_row;
    const int64_t c = i - t * chunks_per_row;
    const int64_t slot = load_index(a.loc, t, a.loc_is_int64);
    const int4 lo = ld_nc_v4(src + t * src_stride + (c << 5));
    const int4 hi = ld_nc_v4(src + t * src_stride + (c << 5) + 16);
    const uint32_t in[8] = {(uint32_t)lo.x, (uint32_t)lo.y, (uint32_t)lo.z, (uint32_t)lo.w,
                            (uint32_t)hi.x, (uint32_t)hi.y, (uint32_t)hi.z, (uint32_t)hi.w};
    uint32_t out[4];
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const float2 p0 = unpack2<T>(in[2 * w]), p1 = unpack2<T>(in[2 * w + 1]);
      const uint32_t w0 = in[2 * w], w1 = in[2 * w + 1];
      out[w] = to_e4m3_torch<T>(p0.x, w0 >> 8 & 0x80u, scale, has_scale) |
               to_e4m3_torch<T>(p0.y, w0 >> 24 & 0x80u, scale, has_scale) << 8 |
               to_e4m3_torch<T>(p1.x, w1 >> 8 & 0x80u, scale, has_scale) << 16 |
               to_e4m3_torch<T>(p1.y, w1 >> 24 & 0x80u, scale, has_scale) << 24;
    }
    *reinterpret_cast<int4*>(cache + slot * cache_stride + (c << 4)) =
        make_int4((int)out[0], (int)out[1], (int)out[2], (int)out[3]);
  }
}

// FP8 KV cache, prefill side: up-cast the cached prefix of every request of an extend batch into a compact bf16/f16
// staging pool, prefix token `pos` of request b -> row offsets[b] + pos, so that the tcgen05 extend kernel (which wants
// 16-bit K-major tiles) reads the prefix from there and the new tokens from the step's activations.  Prefill is
// tensor-bound: the staging traffic (prefix_tokens * Hkv * D * (1 + 2 + 2) B per layer, 31 MB for a 6k prefix of
// Llama-8B) is ~5 us of HBM time next to ~1 ms of attention math.  Each thread turns one 16-byte fp8 chunk into two
// 16-byte stores; x * scale is rounded once to T (scale = 1: exact).
struct KvGatherArgs {
  uint8_t* out[2];
  const uint8_t* cache[2];
  const int32_t* req_to_token;
  int64_t req_to_token_stride;
  const void* req_pool_indices;
  int index_is_int64;
  const int32_t* prefix_lens;
  const int32_t* offsets;
  int64_t row_elems;        // Hkv * D
  int64_t cache_stride;     // bytes (= elements) per token row of the fp8 cache
  float scale[2];
};

template <typename T>
__global__ void __launch_bounds__(256) kv_gather_dequant_fp8_kernel(const KvGatherArgs a) {
  pdl_trigger();
  pdl_wait();
  const int b = blockIdx.z;
  const int which = blockIdx.y;
  const int prefix = a.prefix_lens[b];
  const int64_t chunks_per_row = a.row_elems >> 4;
  const int64_t total = (int64_t)prefix * chunks_per_row;
  const int64_t req = load_index(a.req_pool_indices, b, a.index_is_int64);
  const int32_t* row = a.req_to_token + req * a.req_to_token_stride;
  const uint8_t* cache = a.cache[which];
  uint8_t* out = a.out[which] + (int64_t)a.offsets[b] * a.row_elems * 2;
  const float scale = a.scale[which];
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t pos = i / chunks_per_row;
    const int64_t c = i - pos * chunks_per_row;
    const int4 v = ld_nc_v4(cache + (int64_t)__ldg(row + pos) * a.cache_stride + (c << 4));
    const uint32_t w[4] = {(uint32_t)v.x, (uint32_t)v.y, (uint32_t)v.z, (uint32_t)v.w};
    uint32_t r[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      uint32_t h2;
      const uint16_t pair = (uint16_t)(w[j >> 1] >> ((j & 1) * 16));
      asm("cvt.rn.f16x2.e4m3x2 %0, %1;\n" : "=r"(h2) : "h"(pair));
      const float2 f = __half22float2(*reinterpret_cast<__half2*>(&h2));
      r[j] = pack2<T>(f.x * scale, f.y * scale);
    }
    int4* dst = reinterpret_cast<int4*>(out + (pos * a.row_elems + (c << 4)) * 2);
    dst[0] = make_int4((int)r[0], (int)r[1], (int)r[2], (int)r[3]);
    dst[1] = make_int4((int)r[4], (int)r[5], (int)r[6], (int)r[7]);
  }
}

}  // namespace b200

B200_API int b200_set_kv_buffer(void* k_cache, void* v_cache, const void* k, const void* v,
                                  const void* loc, int loc_is_int64, int64_t num_tokens,
                                  int64_t k_row_elems, int64_t v_row_elems,
                                  int64_t k_token_stride, int64_t v_token_stride,
                                  int64_t k_cache_token_stride, int64_t v_cache_token_stride,
                                  int elem_bytes, void* stream) {
  using namespace b200;
  B200_REQUIRE(num_tokens >= 0, "set_kv_buffer: negative num_tokens");
  if (num_tokens == 0) return B200_OK;
  B200_REQUIRE(k_cache && v_cache && k && v && loc, "set_kv_buffer: null pointer");
  B200_REQUIRE(elem_bytes == 1 || elem_bytes == 2 || elem_bytes == 4, "set_kv_buffer: elem_bytes %d", elem_bytes);
  B200_REQUIRE((k_row_elems * elem_bytes) % 16 == 0 && (v_row_elems * elem_bytes) % 16 == 0,
               "set_kv_buffer: row bytes must be a multiple of 16 (k %lld, v %lld elems x %d B)",
               (long long)k_row_elems, (long long)v_row_elems, elem_bytes);
  B200_REQUIRE((k_token_stride * elem_bytes) % 16 == 0 && (v_token_stride * elem_bytes) % 16 == 0 &&
                   (k_cache_token_stride * elem_bytes) % 16 == 0 &&
                   (v_cache_token_stride * elem_bytes) % 16 == 0,
               "set_kv_buffer: token strides must be 16-byte multiples");
  B200_REQUIRE(((uintptr_t)k_cache | (uintptr_t)v_cache | (uintptr_t)k | (uintptr_t)v) % 16 == 0,
               "set_kv_buffer: pointers must be 16-byte aligned");
  KvWriteArgs a;
  a.cache[0] = (uint8_t*)k_cache;
  a.cache[1] = (uint8_t*)v_cache;
  a.src[0] = (const uint8_t*)k;
  a.src[1] = (const uint8_t*)v;
  a.row_bytes[0] = k_row_elems * elem_bytes;
  a.row_bytes[1] = v_row_elems * elem_bytes;
  a.src_stride_bytes[0] = k_token_stride * elem_bytes;
  a.src_stride_bytes[1] = v_token_stride * elem_bytes;
  a.cache_stride_bytes[0] = k_cache_token_stride * elem_bytes;
  a.cache_stride_bytes[1] = v_cache_token_stride * elem_bytes;
  a.loc = loc;
  a.loc_is_int64 = loc_is_int64;
  a.num_tokens = num_tokens;
  const int64_t max_row = a.row_bytes[0] > a.row_bytes[1] ? a.row_bytes[0] : a.row_bytes[1];
  const int64_t total = num_tokens * (max_row >> 4);
  int64_t blocks = ceil_div(total, 256);
  const int64_t cap = (int64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  return launch_k(kv_write_kernel, dim3((unsigned)blocks, 2), dim3(256), 0, (cudaStream_t)stream, 1, true, a);
}

B200_API int b200_set_kv_buffer_fp8(void* k_cache, void* v_cache, const void* k, const void* v, const void* loc,
                                    int loc_is_int64, int64_t num_tokens, int64_t k_row_elems, int64_t v_row_elems,
                                    int64_t k_token_stride, int64_t v_token_stride, int64_t k_cache_token_stride,
                                    int64_t v_cache_token_stride, float k_scale, float v_scale, int src_dtype,
                                    void* stream) {
  using namespace b200;
  B200_REQUIRE(num_tokens >= 0, "set_kv_buffer_fp8: negative num_tokens");
  if (num_tokens == 0) return B200_OK;
  B200_REQUIRE(k_cache && v_cache && k && v && loc, "set_kv_buffer_fp8: null pointer");
  B200_REQUIRE(k_row_elems % 16 == 0 && v_row_elems % 16 == 0, "set_kv_buffer_fp8: row elements must be a multiple of 16");
  B200_REQUIRE(k_token_stride % 8 == 0 && v_token_stride % 8 == 0 && k_cache_token_stride % 16 == 0 &&
                   v_cache_token_stride % 16 == 0,
               "set_kv_buffer_fp8: token strides must be 16-byte multiples");
  B200_REQUIRE(((uintptr_t)k_cache | (uintptr_t)v_cache | (uintptr_t)k | (uintptr_t)v) % 16 == 0,
               "set_kv_buffer_fp8: pointers must be 16-byte aligned");
  KvWriteArgs a;
  a.cache[0] = (uint8_t*)k_cache;
  a.cache[1] = (uint8_t*)v_cache;
  a.src[0] = (const uint8_t*)k;
  a.src[1] = (const uint8_t*)v;
  a.row_bytes[0] = k_row_elems;
  a.row_bytes[1] = v_row_elems;
  a.src_stride_bytes[0] = k_token_stride * 2;
  a.src_stride_bytes[1] = v_token_stride * 2;
  a.cache_stride_bytes[0] = k_cache_token_stride;
  a.cache_stride_bytes[1] = v_cache_token_stride;
  a.loc = loc;
  a.loc_is_int64 = loc_is_int64;
  a.num_tokens = num_tokens;
  const int64_t max_row = k_row_elems > v_row_elems ? k_row_elems : v_row_elems;
  int64_t blocks = ceil_div(num_tokens * (max_row >> 4), 256);
  const int64_t cap = (int64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  const dim3 grid((unsigned)blocks, 2);
  if (src_dtype == B200_BF16)
    return launch_k(kv_write_fp8_kernel<__nv_bfloat16>, grid, dim3(256), 0, (cudaStream_t)stream, 1, true, a, k_scale,
                    v_scale);
  if (src_dtype == B200_F16)
    return launch_k(kv_write_fp8_kernel<__half>, grid, dim3(256), 0, (cudaStream_t)stream, 1, true, a, k_scale, v_scale);
  return fail(B200_ERR_UNSUPPORTED, "set_kv_buffer_fp8: source dtype %d not supported (bf16/f16)", src_dtype);
}

B200_API int b200_fp8_kv_gather_dequant(void* k_out, void* v_out, const void* k_cache, const void* v_cache,
                                        const int32_t* req_to_token, int64_t req_to_token_stride,
                                        const void* req_pool_indices, int index_is_int64, const int32_t* prefix_lens,
                                        const int32_t* offsets, int batch, int max_prefix_len, int64_t row_elems,
                                        int64_t cache_token_stride, float k_scale, float v_scale, int out_dtype,
                                        void* stream) {
  using namespace b200;
  B200_REQUIRE(batch >= 0 && max_prefix_len >= 0, "fp8_kv_gather_dequant: negative sizes");
  if (batch == 0 || max_prefix_len == 0) return B200_OK;
  B200_REQUIRE(k_out && v_out && k_cache && v_cache && req_to_token && req_pool_indices && prefix_lens && offsets,
               "fp8_kv_gather_dequant: null pointer");
  B200_REQUIRE(row_elems % 16 == 0 && cache_token_stride % 16 == 0, "fp8_kv_gather_dequant: rows must be 16-byte multiples");
  B200_REQUIRE(((uintptr_t)k_out | (uintptr_t)v_out | (uintptr_t)k_cache | (uintptr_t)v_cache) % 16 == 0,
               "fp8_kv_gather_dequant: pointers must be 16-byte aligned");
  B200_REQUIRE(batch <= 65535, "fp8_kv_gather_dequant: batch %d exceeds grid.z", batch);
  B200_REQUIRE(k_scale > 0.f && v_scale > 0.f, "fp8_kv_gather_dequant: scales must be positive");
  KvGatherArgs a;
  a.out[0] = (uint8_t*)k_out; a.out[1] = (uint8_t*)v_out;
  a.cache[0] = (const uint8_t*)k_cache; a.cache[1] = (const uint8_t*)v_cache;
  a.req_to_token = req_to_token; a.req_to_token_stride = req_to_token_stride;
  a.req_pool_indices = req_pool_indices; a.index_is_int64 = index_is_int64;
  a.prefix_lens = prefix_lens; a.offsets = offsets;
  a.row_elems = row_elems; a.cache_stride = cache_token_stride;
  a.scale[0] = k_scale; a.scale[1] = v_scale;
  int64_t blocks = ceil_div((int64_t)max_prefix_len * (row_elems >> 4), 256);
  const int64_t cap = ceil_div((int64_t)sm_count() * 8, batch);
  if (blocks > cap) blocks = cap;
  const dim3 grid((unsigned)blocks, 2, (unsigned)batch);
  if (out_dtype == B200_BF16)
    return launch_k(kv_gather_dequant_fp8_kernel<__nv_bfloat16>, grid, dim3(256), 0, (cudaStream_t)stream, 1, true, a);
  if (out_dtype == B200_F16)
    return launch_k(kv_gather_dequant_fp8_kernel<__half>, grid, dim3(256), 0, (cudaStream_t)stream, 1, true, a);
  return fail(B200_ERR_UNSUPPORTED, "fp8_kv_gather_dequant: output dtype %d not supported (bf16/f16)", out_dtype);
}
