// Element-wise steps either side of the hot path: fused add + RMSNorm,
// SiLU-and-mul, NeoX RoPE with a cos/sin cache.  Reference: sgl-kernel/csrc/elementwise/
// {fused_add_rms_norm_kernel,activation,rope}.cu; numerics follow the layers' forward_native
// (python/sglang/srt/layers/layernorm.py:135-172, activation.py:60-62, rotary_embedding.py:50-75)
// evaluated in fp32.  HBM-bound: 16-byte accesses, one row per CTA.
#include <mutex>

#include "common.cuh"

namespace b200 {

template <typename T>
__global__ void __launch_bounds__(256) rmsnorm_kernel(T* __restrict__ out, T* __restrict__ residual,
                                                      const T* __restrict__ x, const T* __restrict__ w,
                                                      int64_t H, float eps) {
  extern __shared__ float row_f[];  // H floats
  __shared__ float red[8];
  pdl_trigger();
  pdl_wait();
  const int64_t row = blockIdx.x;
  const T* xr = x + row * H;
  T* rr = residual ? residual + row * H : nullptr;
  float ss = 0.f;
  for (int64_t c = threadIdx.x * 8; c < H; c += 256 * 8) {
    const int4 xv = *reinterpret_cast<const int4*>(xr + c);
    const uint32_t* xu = reinterpret_cast<const uint32_t*>(&xv);
    float f[8];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 a = unpack2<T>(xu[i]);
      f[2 * i] = a.x;
      f[2 * i + 1] = a.y;
    }
    if (rr) {
      const int4 rv = *reinterpret_cast<const int4*>(rr + c);
      const uint32_t* ru = reinterpret_cast<const uint32_t*>(&rv);
      int4 nv;
      uint32_t* nu = reinterpret_cast<uint32_t*>(&nv);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 b = unpack2<T>(ru[i]);
        f[2 * i] += b.x;
        f[2 * i + 1] += b.y;
        nu[i] = pack2<T>(f[2 * i], f[2 * i + 1]);
      }
      *reinterpret_cast<int4*>(rr + c) = nv;  // residual = (x + residual) rounded to T
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      row_f[c + i] = f[i];
      ss += f[i] * f[i];
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  float tot = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) tot += red[i];
  const float rstd = rsqrtf(tot / (float)H + eps);
  T* orow = out + row * H;
  for (int64_t c = threadIdx.x * 8; c < H; c += 256 * 8) {
    const int4 wv = *reinterpret_cast<const int4*>(w + c);
    const uint32_t* wu = reinterpret_cast<const uint32_t*>(&wv);
    int4 ov;
    uint32_t* ou = reinterpret_cast<uint32_t*>(&ov);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 ww = unpack2<T>(wu[i]);
      ou[i] = pack2<T>(row_f[c + 2 * i] * rstd * ww.x, row_f[c + 2 * i + 1] * rstd * ww.y);
    }
    *reinterpret_cast<int4*>(orow + c) = ov;
  }
}

template <typename T>
__global__ void __launch_bounds__(256) silu_and_mul_kernel(T* __restrict__ out, const T* __restrict__ x,
                                                           int64_t M, int64_t H) {
  pdl_trigger();
  pdl_wait();
  const int64_t packs = H / 8;
  const int64_t total = M * packs;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = i / packs;
    const int64_t c = (i - row * packs) * 8;
    const int4 gv = *reinterpret_cast<const int4*>(x + row * 2 * H + c);
    const int4 uv = *reinterpret_cast<const int4*>(x + row * 2 * H + H + c);
    const uint32_t* gu = reinterpret_cast<const uint32_t*>(&gv);
    const uint32_t* uu = reinterpret_cast<const uint32_t*>(&uv);
    int4 ov;
    uint32_t* ou = reinterpret_cast<uint32_t*>(&ov);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float2 g2 = unpack2<T>(gu[k]);
      const float2 u2 = unpack2<T>(uu[k]);
      // F.silu(gate) is rounded to T before the multiply in forward_native
      const float s0 = DT<T>::to_f(DT<T>::from_f(g2.x / (1.f + __expf(-g2.x))));
      const float s1 = DT<T>::to_f(DT<T>::from_f(g2.y / (1.f + __expf(-g2.y))));
      ou[k] = pack2<T>(s0 * u2.x, s1 * u2.y);
    }
    *reinterpret_cast<int4*>(out + row * H + c) = ov;
  }
}

// NeoX rotary embedding, in place on q and k.  One CTA per token; thread -> (head, pair index).
template <typename T>
__global__ void __launch_bounds__(256) rope_neox_kernel(T* __restrict__ q, T* __restrict__ k,
                                                        const float* __restrict__ cos_sin, const int64_t* __restrict__ pos,
                                                        int num_q_heads, int num_kv_heads, int head_dim, int rot_dim,
                                                        int64_t q_stride, int64_t k_stride) {
  pdl_trigger();
  pdl_wait();
  const int64_t tok = blockIdx.x;
  const int half = rot_dim / 2;
  const float* cs = cos_sin + pos[tok] * rot_dim;
  const int total = (num_q_heads + num_kv_heads) * half;
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const int h = i / half;
    const int j = i - h * half;
    T* base = h < num_q_heads ? q + tok * q_stride + (int64_t)h * head_dim
                              : k + tok * k_stride + (int64_t)(h - num_q_heads) * head_dim;
    const float c = cs[j], s = cs[half + j];
    const float x1 = DT<T>::to_f(base[j]);
    const float x2 = DT<T>::to_f(base[half + j]);
    base[j] = DT<T>::from_f(x1 * c - x2 * s);
    base[half + j] = DT<T>::from_f(x2 * c + x1 * s);
  }
}

}  // namespace b200

using namespace b200;

// ---- greedy sampling: row-wise argmax of the logits (sampler.py greedy path: torch.argmax(logits, -1)) -----------------
// A row is split over `parts` CTAs (grid.y) so that 64 decode rows still cover every SM; each CTA reduces its
// slice with 16-byte loads and publishes one 64-bit key (order-preserving float bits << 32 | ~index) with
// atomicMax, so equal maxima resolve to the LOWEST index whatever the arrival order; the last CTA of a row
// (counter) decodes the key into the int64 token id and resets key and counter (self-cleaning scratch).
// key_out != nullptr (vocab-parallel greedy sampling): instead of the token id the row's key is stored with the index
// made global (index_offset = first vocabulary row of this TP rank) and one float bit dropped so that it is a
// NON-NEGATIVE int64 — ranks then combine their keys with one 8-byte-per-row max all-reduce and decode the winner
// with argmax_key_decode_kernel.  The dropped bit is always zero for bf16 / f16 logits widened to fp32.
__device__ __forceinline__ long long signed_safe_key(unsigned long long key, uint32_t index_offset) {
  const uint32_t idx = 0xFFFFFFFFu - (uint32_t)(key & 0xFFFFFFFFull) + index_offset;
  return (long long)(((key >> 33) << 32) | (unsigned long long)(0xFFFFFFFFu - idx));
}

__global__ void argmax_key_decode_kernel(int64_t* __restrict__ out, const long long* __restrict__ keys, int64_t rows) {
  pdl_trigger();
  pdl_wait();
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < rows) out[i] = (int64_t)(0xFFFFFFFFu - (uint32_t)((unsigned long long)keys[i] & 0xFFFFFFFFull));
}

template <typename T>
__global__ void __launch_bounds__(256) argmax_kernel(int64_t* __restrict__ out, const T* __restrict__ x, int64_t V,
                                                     int64_t row_stride, unsigned long long* __restrict__ keys,
                                                     unsigned int* __restrict__ counters, long long* __restrict__ key_out,
                                                     uint32_t index_offset) {
  __shared__ unsigned long long s_key[8];
  pdl_trigger();
  pdl_wait();
  const int64_t row = blockIdx.x;
  const int parts = gridDim.y;
  const int64_t per = ((V + parts - 1) / parts + 7) & ~(int64_t)7;
  const int64_t lo = (int64_t)blockIdx.y * per;
  const int64_t hi = lo + per < V ? lo + per : V;
  const T* xr = x + row * row_stride;
  float best = -INFINITY;
  int64_t best_i = V;  // NaN-free logits assumed (as torch's fast path); V = "none seen"
  for (int64_t c = lo + threadIdx.x * 8; c < hi; c += 256 * 8) {
    if (c + 8 <= hi) {
      const int4 v = ld_nc_v4(xr + c);
      const uint32_t* u = reinterpret_cast<const uint32_t*>(&v);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 f = unpack2<T>(u[i]);
        if (f.x > best) { best = f.x; best_i = c + 2 * i; }
        if (f.y > best) { best = f.y; best_i = c + 2 * i + 1; }
      }
    } else {
      for (int64_t j = c; j < hi; ++j) {
        const float f = DT<T>::to_f(xr[j]);
        if (f > best) { best = f; best_i = j; }
      }
    }
  }
  // order-preserving key: larger value wins, then smaller index
  uint32_t fb = __float_as_uint(best);
  fb = (fb & 0x80000000u) ? ~fb : (fb | 0x80000000u);
  unsigned long long key = ((unsigned long long)fb << 32) | (unsigned long long)(0xFFFFFFFFu - (uint32_t)best_i);
  if (best_i >= V) key = 0ull;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
    key = other > key ? other : key;
  }
  if ((threadIdx.x & 31) == 0) s_key[threadIdx.x >> 5] = key;
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int w = 1; w < 8; ++w) key = s_key[w] > key ? s_key[w] : key;
    if (parts == 1) {
      if (key_out != nullptr) key_out[row] = signed_safe_key(key, index_offset);
      else out[row] = (int64_t)(0xFFFFFFFFu - (uint32_t)(key & 0xFFFFFFFFull));
    } else {
      atomicMax(keys + row, key);
      __threadfence();
      if (atomicAdd(counters + row, 1u) == (unsigned)parts - 1u) {
        __threadfence();
        const unsigned long long k = atomicExch(keys + row, 0ull);
        counters[row] = 0u;
        if (key_out != nullptr) key_out[row] = signed_safe_key(k, index_offset);
        else out[row] = (int64_t)(0xFFFFFFFFu - (uint32_t)(k & 0xFFFFFFFFull));
      }
    }
  }
}

B200_API int b200_rmsnorm(void* out, void* residual, const void* x, const void* weight, int64_t M, int64_t H,
                            float eps, int dtype, void* stream) {
  B200_REQUIRE(M >= 0 && H > 0, "rmsnorm: bad sizes");
  if (M == 0) return B200_OK;
  B200_REQUIRE(out && x && weight, "rmsnorm: null pointer");
  B200_REQUIRE(H % 8 == 0 && H <= 16384, "rmsnorm: H %lld must be a multiple of 8 and <= 16384", (long long)H);
  cudaStream_t st = (cudaStream_t)stream;
  const size_t smem = (size_t)H * sizeof(float);
  if (dtype == B200_BF16) {
    auto kern = rmsnorm_kernel<__nv_bfloat16>;
    if (smem > 48 * 1024) B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    return launch_k(kern, dim3((unsigned)M), dim3(256), smem, st, 1, true, (__nv_bfloat16*)out, (__nv_bfloat16*)residual,
                    (const __nv_bfloat16*)x, (const __nv_bfloat16*)weight, H, eps);
  } else if (dtype == B200_F16) {
    auto kern = rmsnorm_kernel<__half>;
    if (smem > 48 * 1024) B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    return launch_k(kern, dim3((unsigned)M), dim3(256), smem, st, 1, true, (__half*)out, (__half*)residual, (const __half*)x,
                    (const __half*)weight, H, eps);
  } else {
    return fail(B200_ERR_UNSUPPORTED, "rmsnorm: dtype %d not supported", dtype);
  }
  B200_LAUNCH_OK();
  return B200_OK;
}

B200_API int b200_silu_and_mul(void* out, const void* x, int64_t M, int64_t H, int dtype, void* stream) {
  B200_REQUIRE(M >= 0 && H > 0, "silu_and_mul: bad sizes");
  if (M == 0) return B200_OK;
  B200_REQUIRE(out && x, "silu_and_mul: null pointer");
  B200_REQUIRE(H % 8 == 0, "silu_and_mul: H must be a multiple of 8");
  cudaStream_t st = (cudaStream_t)stream;
  int64_t blocks = ceil_div(M * (H / 8), 256);
  const int64_t cap = (int64_t)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  if (dtype == B200_BF16)
    return launch_k(silu_and_mul_kernel<__nv_bfloat16>, dim3((unsigned)blocks), dim3(256), 0, st, 1, true, (__nv_bfloat16*)out,
                    (const __nv_bfloat16*)x, M, H);
  if (dtype == B200_F16)
    return launch_k(silu_and_mul_kernel<__half>, dim3((unsigned)blocks), dim3(256), 0, st, 1, true, (__half*)out,
                    (const __half*)x, M, H);
  return fail(B200_ERR_UNSUPPORTED, "silu_and_mul: dtype %d not supported", dtype);
}

B200_API int b200_rope_neox(void* q, void* k, const float* cos_sin_cache, const int64_t* positions,
                              int64_t num_tokens, int num_q_heads, int num_kv_heads, int head_dim, int rot_dim,
                              int64_t q_token_stride, int64_t k_token_stride, int dtype, void* stream) {
  B200_REQUIRE(num_tokens >= 0, "rope_neox: bad sizes");
  if (num_tokens == 0) return B200_OK;
  B200_REQUIRE(q && k && cos_sin_cache && positions, "rope_neox: null pointer");
  B200_REQUIRE(rot_dim % 2 == 0 && rot_dim <= head_dim, "rope_neox: rot_dim %d", rot_dim);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == B200_BF16)
    return launch_k(rope_neox_kernel<__nv_bfloat16>, dim3((unsigned)num_tokens), dim3(256), 0, st, 1, true, (__nv_bfloat16*)q,
                    (__nv_bfloat16*)k, cos_sin_cache, positions, num_q_heads, num_kv_heads, head_dim, rot_dim, q_token_stride,
                    k_token_stride);
  if (dtype == B200_F16)
    return launch_k(rope_neox_kernel<__half>, dim3((unsigned)num_tokens), dim3(256), 0, st, 1, true, (__half*)q, (__half*)k,
                    cos_sin_cache, positions, num_q_heads, num_kv_heads, head_dim, rot_dim, q_token_stride, k_token_stride);
  return fail(B200_ERR_UNSUPPORTED, "rope_neox: dtype %d not supported", dtype);
}

B200_API int b200_argmax_workspace_bytes(int64_t rows, size_t* bytes) {
  B200_REQUIRE(bytes != nullptr && rows >= 0, "argmax_workspace_bytes: bad arguments");
  *bytes = (size_t)rows * 16;  // one u64 key + one u32 counter (padded) per row; zero-initialised once, self-cleaning
  return B200_OK;
}

static int run_argmax(int64_t* out, long long* key_out, uint32_t index_offset, const void* x, int64_t rows, int64_t V,
                      int64_t row_stride, int dtype, void* workspace, size_t workspace_bytes, void* stream) {
  B200_REQUIRE(rows >= 0 && V > 0 && V < (1ll << 32), "argmax: bad sizes");
  if (rows == 0) return B200_OK;
  B200_REQUIRE((out || key_out) && x, "argmax: null pointer");
  B200_REQUIRE(row_stride % 8 == 0 && ((uintptr_t)x % 16) == 0, "argmax: rows must be 16-byte aligned");
  B200_REQUIRE(rows <= 65535 * 32768ll, "argmax: too many rows");
  int parts = (int)((2 * (int64_t)sm_count() + rows - 1) / rows);  // ~2 CTAs per SM
  const int64_t max_parts = ceil_div(V, 2048);
  if (parts > max_parts) parts = (int)max_parts;
  if (parts < 1) parts = 1;
  if (workspace == nullptr || workspace_bytes < (size_t)rows * 16) parts = 1;
  unsigned long long* keys = reinterpret_cast<unsigned long long*>(workspace);
  unsigned int* counters = parts > 1 ? reinterpret_cast<unsigned int*>(keys + rows) : nullptr;
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid((unsigned)rows, (unsigned)parts);
  if (dtype == B200_BF16)
    return launch_k(argmax_kernel<__nv_bfloat16>, grid, dim3(256), 0, st, 1, true, out, (const __nv_bfloat16*)x, V,
                    row_stride, keys, counters, key_out, index_offset);
  if (dtype == B200_F16)
    return launch_k(argmax_kernel<__half>, grid, dim3(256), 0, st, 1, true, out, (const __half*)x, V, row_stride, keys,
                    counters, key_out, index_offset);
  return fail(B200_ERR_UNSUPPORTED, "argmax: dtype %d not supported (bf16/f16)", dtype);
}

B200_API int b200_argmax(int64_t* out, const void* x, int64_t rows, int64_t V, int64_t row_stride, int dtype,
                         void* workspace, size_t workspace_bytes, void* stream) {
  return run_argmax(out, nullptr, 0u, x, rows, V, row_stride, dtype, workspace, workspace_bytes, stream);
}

B200_API int b200_argmax_key(int64_t* key_out, const void* x, int64_t rows, int64_t V, int64_t row_stride, int dtype,
                             int64_t index_offset, void* workspace, size_t workspace_bytes, void* stream) {
  B200_REQUIRE(index_offset >= 0 && index_offset + V < (1ll << 32), "argmax_key: index offset out of range");
  return run_argmax(nullptr, reinterpret_cast<long long*>(key_out), (uint32_t)index_offset, x, rows, V, row_stride, dtype,
                    workspace, workspace_bytes, stream);
}

B200_API int b200_argmax_key_decode(int64_t* out, const int64_t* keys, int64_t rows, void* stream) {
  B200_REQUIRE(rows >= 0, "argmax_key_decode: bad sizes");
  if (rows == 0) return B200_OK;
  B200_REQUIRE(out && keys, "argmax_key_decode: null pointer");
  return launch_k(argmax_key_decode_kernel, dim3((unsigned)ceil_div(rows, 128)), dim3(128), 0, (cudaStream_t)stream, 1, true,
                  out, reinterpret_cast<const long long*>(keys), rows);
}

// ---- greedy argmax without any allocation: scratch in a file-scope device array ---------------------------------------------
// For callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
// The scratch of b200_argmax (one u64 key + one u32 counter, padded to 16 bytes, per row) lives in g_argmax_scratch:
// 4096 rows x 16 B = 64 KB, held by every GPU that loads the library until the process exits.  In host code the array's
// name is only a host-side stand-in, so the device address is looked up ONCE per device with cudaGetSymbolAddress (with
// that device current) and that pointer is what the kernel gets.  Every call on a device shares the array: it is cleared
// on the caller's stream before each launch (an earlier call that was aborted may have left keys behind), and two calls
// that overlap on the GPU (different streams, no dependency) race on it.
constexpr int64_t kArgmaxStaticMaxRows = 4096;
static __device__ __align__(256) unsigned char g_argmax_scratch[kArgmaxStaticMaxRows * 16];

static int argmax_static_scratch(void** ptr) {
  constexpr int kMaxDevices = 64;
  static std::mutex mu;
  static void* addr[kMaxDevices] = {};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < kMaxDevices, "argmax_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (addr[dev] == nullptr) {
    void* p = nullptr;
    B200_CUDA_OK(cudaGetSymbolAddress(&p, g_argmax_scratch));  // this device's instance of the array
    B200_REQUIRE(p != nullptr && ((uintptr_t)p % 256) == 0, "argmax_static: bad scratch address");
    addr[dev] = p;
  }
  *ptr = addr[dev];
  return B200_OK;
}

B200_API int b200_argmax_static(int64_t* out, const void* x, int64_t rows, int64_t V, int64_t row_stride, int dtype,
                                void* stream) {
  B200_REQUIRE(rows >= 0, "argmax_static: bad sizes");
  if (rows > kArgmaxStaticMaxRows)
    return fail(B200_ERR_WORKSPACE, "argmax_static: %lld rows exceed the built-in scratch of %lld rows (use b200_argmax "
                "with a caller-owned workspace)", (long long)rows, (long long)kArgmaxStaticMaxRows);
  if (rows == 0) return B200_OK;
  void* scratch = nullptr;
  int rc = argmax_static_scratch(&scratch);
  if (rc != B200_OK) return rc;
  const size_t bytes = (size_t)rows * 16;
  B200_CUDA_OK(cudaMemsetAsync(scratch, 0, bytes, (cudaStream_t)stream));  // keys and counters must start at zero
  return run_argmax(out, nullptr, 0u, x, rows, V, row_stride, dtype, scratch, bytes, stream);
}

// ---- decode prologue: token embedding gather (+ positions = seq_lens - 1) -------------------------------------------------
// hidden[t, :] = table[ids[t], :] (VocabParallelEmbedding.forward with a replicated table, vocab_parallel_embedding.py) and,
// when seq_lens is given, positions[t] = seq_lens[t] - 1 (forward_batch_info.py:960-962) — one launch in place of an
// indexing kernel and a subtraction.
__global__ void __launch_bounds__(128) embedding_kernel(uint8_t* __restrict__ out, const uint8_t* __restrict__ table,
                                                        const int64_t* __restrict__ ids, int64_t row_bytes, int64_t vocab,
                                                        int64_t* __restrict__ positions, const int64_t* __restrict__ seq_lens) {
  pdl_trigger();
  pdl_wait();
  const int64_t t = blockIdx.x;
  int64_t id = ids[t];
  if (id < 0 || id >= vocab) id = 0;  // out-of-range ids read row 0 instead of faulting (padded graph slots)
  const int4* src = reinterpret_cast<const int4*>(table + id * row_bytes);
  int4* dst = reinterpret_cast<int4*>(out + t * row_bytes);
  for (int64_t i = threadIdx.x; i < row_bytes / 16; i += blockDim.x) dst[i] = ld_nc_v4(src + i);
  if (positions != nullptr && threadIdx.x == 0) positions[t] = seq_lens[t] - 1;
}

B200_API int b200_embedding(void* out, const void* table, const int64_t* ids, int64_t T, int64_t H, int64_t vocab, int dtype,
                            int64_t* positions, const int64_t* seq_lens, void* stream) {
  B200_REQUIRE(T >= 0 && H > 0 && vocab > 0, "embedding: bad sizes");
  if (T == 0) return B200_OK;
  B200_REQUIRE(out && table && ids, "embedding: null pointer");
  B200_REQUIRE((positions == nullptr) == (seq_lens == nullptr), "embedding: positions and seq_lens go together");
  if (dtype != B200_BF16 && dtype != B200_F16) return fail(B200_ERR_UNSUPPORTED, "embedding: dtype %d not supported", dtype);
  B200_REQUIRE((H * 2) % 16 == 0 && ((uintptr_t)out % 16) == 0 && ((uintptr_t)table % 16) == 0, "embedding: rows must be 16-byte multiples");
  return launch_k(embedding_kernel, dim3((unsigned)T), dim3(128), 0, (cudaStream_t)stream, 1, true, (uint8_t*)out,
                  (const uint8_t*)table, ids, H * 2, vocab, positions, seq_lens);
}
