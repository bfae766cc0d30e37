// Caller-side page-table kernels: the integer work that feeds the hot path every step.
//
// Replaces the reference's Triton kernels (python/sglang/srt/...):
//   alloc_extend_kernel                mem_cache/allocator.py:276-360   (+ alloc_extend_native :583-660, the torch restatement)
//   alloc_decode_kernel                mem_cache/allocator.py:364-401   (+ alloc_decode_kernel_ascend :573-582)
//   write_req_to_token_pool_triton     managers/schedule_batch.py:1920-1953
//   get_last_loc_kernel / _torch       managers/schedule_batch.py:1972-2027
//   compute_position_kernel / _torch   model_executor/forward_batch_info.py:886-957
//   create_flashinfer_kv_indices_triton layers/attention/utils.py:11-47
// All outputs are bit-exact integers.  One CTA per request; the per-request exclusive prefix sums the Triton kernels
// obtain by re-loading every earlier request in every program are block reductions here.  Two fusions the reference
// leaves as TODO ("fuse last_loc into the kernel", allocator.py:411): b200_alloc_extend / b200_alloc_decode can read
// last_loc straight from req_to_token and write the new slots back into req_to_token in the same launch, so one
// launch replaces get_last_loc + alloc_* + write_req_to_token_pool.
#include "common.cuh"

namespace b200 {

constexpr int kPtThreads = 256;

// sum of f(i) for i in [0, n) over the CTA; every thread gets the result
template <typename F>
__device__ __forceinline__ int64_t block_sum(int n, F f, int64_t* s_red) {
  int64_t v = 0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) v += f(i);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();  // s_red may still be read from a previous call
  if (lane == 0) s_red[warp] = v;
  __syncthreads();
  int64_t t = 0;
  for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += s_red[w];
  return t;
}

struct AllocArgs {
  const void* pre_lens;   // [bs] (extend) or null (decode: seq_len - 1)
  const void* seq_lens;   // [bs]
  const void* last_loc;   // [bs] or null -> read req_to_token[req, pre_len - 1] (-1 when pre_len == 0)
  const int64_t* free_pages;
  int64_t* out_indices;   // [sum extend] (extend) / [bs] (decode)
  int64_t* ret_values;    // extend: (new_pages << 32) | sum_extend ; decode: new_pages
  int32_t* req_to_token;  // optional (fusions)
  const void* req_pool_indices;
  int64_t r2t_stride;
  int bs, page_size;
  int lens_is64, pool_is64;
  int write_back;         // store the new slots into req_to_token[req, pre_len : seq_len]
};

__device__ __forceinline__ int64_t pages_of(int64_t len, int page) { return (len + page - 1) / page; }

template <bool DECODE>
__global__ void __launch_bounds__(kPtThreads) alloc_kernel(const AllocArgs a) {
  __shared__ int64_t s_red[kPtThreads / 32];
  const int pid = blockIdx.x;
  const int P = a.page_size;
  pdl_trigger();
  pdl_wait();
  auto seq_of = [&](int i) { return load_index(a.seq_lens, i, a.lens_is64); };
  auto pre_of = [&](int i) { return DECODE ? seq_of(i) - 1 : load_index(a.pre_lens, i, a.lens_is64); };
  const int64_t seq_len = seq_of(pid);
  const int64_t pre_len = pre_of(pid);
  const int64_t extend_len = seq_len - pre_len;
  // inclusive sums over requests 0..pid (allocator.py:291-312)
  const int64_t sum_extend = DECODE ? (int64_t)(pid + 1) : block_sum(pid + 1, [&](int i) { return seq_of(i) - pre_of(i); }, s_red);
  const int64_t sum_pages = block_sum(pid + 1, [&](int i) { return pages_of(seq_of(i), P) - pages_of(pre_of(i), P); }, s_red);
  const int64_t my_pages = pages_of(seq_len, P) - pages_of(pre_len, P);
  const int64_t out_start = sum_extend - extend_len;
  const int64_t page_start = sum_pages - my_pages;
  if (pid == (int)gridDim.x - 1 && threadIdx.x == 0)
    *a.ret_values = DECODE ? sum_pages : ((sum_pages << 32) | sum_extend);

  int64_t req = 0;
  if (a.req_to_token != nullptr) req = load_index(a.req_pool_indices, pid, a.pool_is64);
  int64_t last_loc;
  if (a.last_loc != nullptr) last_loc = load_index(a.last_loc, pid, a.lens_is64);
  else last_loc = pre_len > 0 ? (int64_t)a.req_to_token[req * a.r2t_stride + pre_len - 1] : -1;
  int32_t* wb = (a.write_back && a.req_to_token != nullptr) ? a.req_to_token + req * a.r2t_stride + pre_len : nullptr;

  // token j of the extension (0 <= j < extend_len) lands in:
  //   part 1 (rest of the old partial page)  j <  n1            -> last_loc + 1 + j
  //   parts 2+3 (new pages, in free_pages order)                -> free_pages[page_start + (j - n1) / P] * P + (j - n1) % P
  const int64_t n1 = min(seq_len, pages_of(pre_len, P) * P) - pre_len;
  for (int64_t j = threadIdx.x; j < extend_len; j += blockDim.x) {
    int64_t slot;
    if (j < n1) {
      slot = last_loc + 1 + j;
    } else {
      const int64_t k = j - n1;
      slot = a.free_pages[page_start + k / P] * P + k % P;
    }
    a.out_indices[out_start + j] = slot;
    if (wb != nullptr) wb[j] = (int32_t)slot;
  }
}

// req_to_token[req, pre_len : seq_len] = out_cache_loc[cumsum(extend_lens)[pid] ...] (schedule_batch.py:1920-1953)
__global__ void __launch_bounds__(kPtThreads)
write_req_to_token_kernel(int32_t* req_to_token, int64_t stride, const void* req_pool_indices, int pool_is64,
                          const void* pre_lens, const void* seq_lens, const void* extend_lens, int lens_is64,
                          const void* out_cache_loc, int loc_is64) {
  __shared__ int64_t s_red[kPtThreads / 32];
  const int pid = blockIdx.x;
  pdl_trigger();
  pdl_wait();
  const int64_t req = load_index(req_pool_indices, pid, pool_is64);
  const int64_t pre = load_index(pre_lens, pid, lens_is64);
  const int64_t n = load_index(seq_lens, pid, lens_is64) - pre;
  const int64_t start = block_sum(pid, [&](int i) { return load_index(extend_lens, i, lens_is64); }, s_red);
  int32_t* dst = req_to_token + req * stride + pre;
  for (int64_t j = threadIdx.x; j < n; j += blockDim.x) dst[j] = (int32_t)load_index(out_cache_loc, start + j, loc_is64);
}

// result[i] = prefix_lens[i] > 0 ? req_to_token[req[i], prefix_lens[i] - 1] : -1 (schedule_batch.py:1972-1982)
__global__ void __launch_bounds__(kPtThreads)
get_last_loc_kernel(void* result, const int32_t* req_to_token, int64_t stride, const void* req_pool_indices, int pool_is64,
                    const void* prefix_lens, int lens_is64, int n) {
  pdl_trigger();
  pdl_wait();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int64_t pl = load_index(prefix_lens, i, lens_is64);
  const int64_t v = pl > 0 ? (int64_t)req_to_token[load_index(req_pool_indices, i, pool_is64) * stride + pl - 1] : -1;
  if (lens_is64) reinterpret_cast<int64_t*>(result)[i] = v;
  else reinterpret_cast<int32_t*>(result)[i] = (int32_t)v;
}

// positions[start_b + j] = prefix_b + j, extend_start_loc[b] = start_b (forward_batch_info.py:913-957)
__global__ void __launch_bounds__(kPtThreads)
compute_position_kernel(int64_t* positions, int32_t* extend_start_loc, const void* prefix_lens, const void* extend_lens,
                        int lens_is64) {
  __shared__ int64_t s_red[kPtThreads / 32];
  const int pid = blockIdx.x;
  pdl_trigger();
  pdl_wait();
  const int64_t prefix = prefix_lens ? load_index(prefix_lens, pid, lens_is64) : 0;
  const int64_t n = load_index(extend_lens, pid, lens_is64);
  const int64_t start = block_sum(pid, [&](int i) { return load_index(extend_lens, i, lens_is64); }, s_red);
  for (int64_t j = threadIdx.x; j < n; j += blockDim.x) positions[start + j] = prefix + j;
  if (threadIdx.x == 0) extend_start_loc[pid] = (int32_t)start;
}

// kv_indices[kv_indptr[b] + j] = req_to_token[req_b, kv_start_b + j], j < len_b (attention/utils.py:11-47)
__global__ void __launch_bounds__(kPtThreads)
kv_indices_kernel(int32_t* kv_indices, const int32_t* req_to_token, int64_t stride, const void* req_pool_indices,
                  int pool_is64, const void* kernel_lens, int lens_is64, const int32_t* kv_indptr,
                  const int32_t* kv_start_idx) {
  const int pid = blockIdx.x;
  pdl_trigger();
  pdl_wait();
  const int64_t req = load_index(req_pool_indices, pid, pool_is64);
  const int64_t start = kv_start_idx ? (int64_t)kv_start_idx[pid] : 0;
  const int64_t n = load_index(kernel_lens, pid, lens_is64);
  const int32_t* src = req_to_token + req * stride + start;
  int32_t* dst = kv_indices + kv_indptr[pid];
  for (int64_t j = threadIdx.x; j < n; j += blockDim.x) dst[j] = src[j];
}

static int run_alloc(bool decode, AllocArgs a, cudaStream_t st) {
  B200_REQUIRE(a.bs >= 0 && a.page_size >= 1, "alloc: bad batch / page size");
  if (a.bs == 0) return B200_OK;
  B200_REQUIRE(a.seq_lens && a.free_pages && a.out_indices && a.ret_values, "alloc: null pointer");
  B200_REQUIRE(decode || a.pre_lens, "alloc_extend: null prefix lens");
  B200_REQUIRE(a.last_loc != nullptr || (a.req_to_token != nullptr && a.req_pool_indices != nullptr),
               "alloc: either last_loc or req_to_token + req_pool_indices must be given");
  B200_REQUIRE(!a.write_back || (a.req_to_token != nullptr && a.req_pool_indices != nullptr),
               "alloc: write_back needs req_to_token + req_pool_indices");
  if (decode) return launch_k(alloc_kernel<true>, dim3(a.bs), dim3(kPtThreads), 0, st, 1, true, a);
  return launch_k(alloc_kernel<false>, dim3(a.bs), dim3(kPtThreads), 0, st, 1, true, a);
}

}  // namespace b200

using namespace b200;

B200_API int b200_alloc_extend(int64_t* out_indices, int64_t* ret_values, const void* prefix_lens, const void* seq_lens,
                               const void* last_loc, int lens_is_int64, const int64_t* free_pages, int batch, int page_size,
                               int32_t* req_to_token, int64_t req_to_token_stride, const void* req_pool_indices,
                               int pool_is_int64, int write_req_to_token, void* stream) {
  AllocArgs a = {};
  a.pre_lens = prefix_lens; a.seq_lens = seq_lens; a.last_loc = last_loc; a.free_pages = free_pages;
  a.out_indices = out_indices; a.ret_values = ret_values; a.req_to_token = req_to_token;
  a.req_pool_indices = req_pool_indices; a.r2t_stride = req_to_token_stride; a.bs = batch; a.page_size = page_size;
  a.lens_is64 = lens_is_int64; a.pool_is64 = pool_is_int64; a.write_back = write_req_to_token;
  return run_alloc(false, a, (cudaStream_t)stream);
}

B200_API int b200_alloc_decode(int64_t* out_indices, int64_t* ret_values, const void* seq_lens, const void* last_loc,
                               int lens_is_int64, const int64_t* free_pages, int batch, int page_size,
                               int32_t* req_to_token, int64_t req_to_token_stride, const void* req_pool_indices,
                               int pool_is_int64, int write_req_to_token, void* stream) {
  AllocArgs a = {};
  a.seq_lens = seq_lens; a.last_loc = last_loc; a.free_pages = free_pages; a.out_indices = out_indices;
  a.ret_values = ret_values; a.req_to_token = req_to_token; a.req_pool_indices = req_pool_indices;
  a.r2t_stride = req_to_token_stride; a.bs = batch; a.page_size = page_size; a.lens_is64 = lens_is_int64;
  a.pool_is64 = pool_is_int64; a.write_back = write_req_to_token;
  return run_alloc(true, a, (cudaStream_t)stream);
}

B200_API int b200_write_req_to_token_pool(int32_t* req_to_token, int64_t req_to_token_stride, const void* req_pool_indices,
                                          int pool_is_int64, const void* prefix_lens, const void* seq_lens,
                                          const void* extend_lens, int lens_is_int64, const void* out_cache_loc,
                                          int loc_is_int64, int batch, void* stream) {
  B200_REQUIRE(batch >= 0, "write_req_to_token_pool: bad batch");
  if (batch == 0) return B200_OK;
  B200_REQUIRE(req_to_token && req_pool_indices && prefix_lens && seq_lens && extend_lens && out_cache_loc,
               "write_req_to_token_pool: null pointer");
  return launch_k(write_req_to_token_kernel, dim3(batch), dim3(kPtThreads), 0, (cudaStream_t)stream, 1, true, req_to_token,
                  req_to_token_stride, req_pool_indices, pool_is_int64, prefix_lens, seq_lens, extend_lens, lens_is_int64,
                  out_cache_loc, loc_is_int64);
}

B200_API int b200_get_last_loc(void* result, const int32_t* req_to_token, int64_t req_to_token_stride,
                               const void* req_pool_indices, int pool_is_int64, const void* prefix_lens, int lens_is_int64,
                               int n, void* stream) {
  B200_REQUIRE(n >= 0, "get_last_loc: bad size");
  if (n == 0) return B200_OK;
  B200_REQUIRE(result && req_to_token && req_pool_indices && prefix_lens, "get_last_loc: null pointer");
  return launch_k(get_last_loc_kernel, dim3((unsigned)ceil_div(n, kPtThreads)), dim3(kPtThreads), 0, (cudaStream_t)stream,
                  1, true, result, req_to_token, req_to_token_stride, req_pool_indices, pool_is_int64, prefix_lens,
                  lens_is_int64, n);
}

B200_API int b200_compute_position(int64_t* positions, int32_t* extend_start_loc, const void* extend_prefix_lens,
                                   const void* extend_seq_lens, int lens_is_int64, int batch, void* stream) {
  B200_REQUIRE(batch >= 0, "compute_position: bad batch");
  if (batch == 0) return B200_OK;
  B200_REQUIRE(positions && extend_start_loc && extend_seq_lens, "compute_position: null pointer");
  return launch_k(compute_position_kernel, dim3(batch), dim3(kPtThreads), 0, (cudaStream_t)stream, 1, true, positions,
                  extend_start_loc, extend_prefix_lens, extend_seq_lens, lens_is_int64);
}

B200_API int b200_create_kv_indices(int32_t* kv_indices, const int32_t* req_to_token, int64_t req_to_token_stride,
                                    const void* req_pool_indices, int pool_is_int64, const void* page_kernel_lens,
                                    int lens_is_int64, const int32_t* kv_indptr, const int32_t* kv_start_idx, int batch,
                                    void* stream) {
  B200_REQUIRE(batch >= 0, "create_kv_indices: bad batch");
  if (batch == 0) return B200_OK;
  B200_REQUIRE(kv_indices && req_to_token && req_pool_indices && page_kernel_lens && kv_indptr,
               "create_kv_indices: null pointer");
  return launch_k(kv_indices_kernel, dim3(batch), dim3(kPtThreads), 0, (cudaStream_t)stream, 1, true, kv_indices,
                  req_to_token, req_to_token_stride, req_pool_indices, pool_is_int64, page_kernel_lens, lens_is_int64,
                  kv_indptr, kv_start_idx);
}
