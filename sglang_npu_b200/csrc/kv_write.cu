// set_kv_buffer: scatter the step's K/V rows into the paged KV pool (bit-exact copy).
// Replaces the two index_put_ launches of MHATokenToKVPool.set_kv_buffer
// (reference python/sglang/srt/mem_cache/memory_pool.py:369-407) with one launch.
//
// HBM-bound byte work: each thread moves one 16-byte chunk of one token row; a warp
// covers 512 contiguous bytes of a row, so both the gather (k/v activations) and the
// scatter (cache rows) are fully coalesced.  blockIdx.y selects K (0) or V (1).
#include "common.cuh"

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
