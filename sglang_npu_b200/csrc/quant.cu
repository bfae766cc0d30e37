// Dynamic activation quantisation feeding the quantised GEMMs.
//   per_token_quant_int8        — reference: python/sglang/srt/layers/quantization/int8_kernel.py:20-83
//                                 (oracle rule: test/srt/cpu/utils.py:24-32, round-half-even)
//   per_token_group_quant_fp8   — reference: sgl-kernel/csrc/gemm/per_token_group_quant_8bit.cu,
//                                 fp8_kernel.py:116-297 (native restatement test_block_fp8.py:24-49)
//   per_token_quant_fp8         — reference: sgl-kernel/csrc/gemm/per_token_quant_fp8.cu
// HBM-bound element-wise work: 16-byte loads, one row (or eight 128-groups) per CTA pass,
// second pass re-reads the row from L1/L2.  All divisions are IEEE (div.rn) and no FMA
// contraction can occur (single multiplies), so the int8 codes are bit-exact vs the oracle.
#include "common.cuh"

namespace b200 {

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

template <int THREADS>
__device__ __forceinline__ float block_max(float v, float* sm) {
  v = warp_max(v);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = sm[0];
#pragma unroll
  for (int i = 1; i < THREADS / 32; ++i) r = fmaxf(r, sm[i]);
  __syncthreads();
  return r;
}

template <typename T>
__device__ __forceinline__ void load8(const T* p, float (&f)[8]) {
  const int4 v = *reinterpret_cast<const int4*>(p);
  const uint32_t* u = reinterpret_cast<const uint32_t*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 x = unpack2<T>(u[i]);
    f[2 * i] = x.x;
    f[2 * i + 1] = x.y;
  }
}

// ---- int8 per token ---------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) per_token_quant_int8_kernel(int8_t* __restrict__ xq, float* __restrict__ scale,
                                                                   const T* __restrict__ x, int64_t K,
                                                                   int64_t x_row_stride) {
  __shared__ float sm[8];
  pdl_trigger();
  pdl_wait();
  const int64_t row = blockIdx.x;
  const T* xr = x + row * x_row_stride;
  float amax = 0.f;
  for (int64_t c = threadIdx.x * 8; c < K; c += 256 * 8) {
    float f[8];
    load8<T>(xr + c, f);
#pragma unroll
    for (int i = 0; i < 8; ++i) amax = fmaxf(amax, fabsf(f[i]));
  }
  amax = fmaxf(block_max<256>(amax, sm), 1e-10f);
  // torch evaluates `127 / absmax` as reciprocal(absmax) * 127 (Tensor.__rtruediv__), i.e. two
  // roundings; the oracle (test/srt/cpu/utils.py:24-32 run by torch) fixes that rule, so do the same.
  const float inv = __fmul_rn(__frcp_rn(amax), 127.f);
  if (threadIdx.x == 0) scale[row] = __fdiv_rn(amax, 127.f);
  int8_t* qr = xq + row * K;
  for (int64_t c = threadIdx.x * 8; c < K; c += 256 * 8) {
    float f[8];
    load8<T>(xr + c, f);
    uint32_t lo = 0, hi = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int a = __float2int_rn(__fmul_rn(f[i], inv));
      const int b = __float2int_rn(__fmul_rn(f[4 + i], inv));
      lo |= (uint32_t)(a & 0xff) << (8 * i);
      hi |= (uint32_t)(b & 0xff) << (8 * i);
    }
    *reinterpret_cast<uint2*>(qr + c) = make_uint2(lo, hi);
  }
}

// ---- fp8 helpers -----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t pack_fp8x4(float a, float b, float c, float d) {
  const __nv_fp8x2_storage_t lo = __nv_cvt_float2_to_fp8x2(make_float2(a, b), __NV_SATFINITE, __NV_E4M3);
  const __nv_fp8x2_storage_t hi = __nv_cvt_float2_to_fp8x2(make_float2(c, d), __NV_SATFINITE, __NV_E4M3);
  return (uint32_t)lo | ((uint32_t)hi << 16);
}

// ---- fp8 per (token, group) -------------------------------------------------------------------
// One warp quantises GROUPS_PER_WARP consecutive groups; a group of G elements is covered by
// G/8 lanes (16-byte load each).
template <typename T>
__global__ void __launch_bounds__(256) per_token_group_quant_fp8_kernel(
    uint8_t* __restrict__ xq, float* __restrict__ scales, const T* __restrict__ x, int64_t M, int64_t K,
    int64_t x_row_stride, int group_size, float eps, int64_t s_row_stride, int64_t s_col_stride) {
  pdl_trigger();
  pdl_wait();  // x is the previous kernel's output (decode: the attention merge, between it and the o_proj GEMM)
  const int lanes_per_group = group_size / 8;       // 16 for G=128
  const int groups_per_warp = 32 / lanes_per_group;
  const int groups_per_row = (int)(K / group_size);
  const int64_t total_groups = M * groups_per_row;
  const int lane = threadIdx.x & 31;
  const int64_t warp_global = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int64_t gidx = warp_global * groups_per_warp + lane / lanes_per_group;
  const int li = lane % lanes_per_group;
  const bool active = gidx < total_groups;
  float f[8];
  float amax = 0.f;
  int64_t row = 0;
  int gcol = 0;
  if (active) {
    row = gidx / groups_per_row;
    gcol = (int)(gidx - row * groups_per_row);
    load8<T>(x + row * x_row_stride + (int64_t)gcol * group_size + li * 8, f);
#pragma unroll
    for (int i = 0; i < 8; ++i) amax = fmaxf(amax, fabsf(f[i]));
  }
  for (int o = lanes_per_group >> 1; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
  if (!active) return;
  amax = fmaxf(amax, eps);
  const float s = __fdiv_rn(amax, 448.f);
  if (li == 0) scales[row * s_row_stride + gcol * s_col_stride] = s;
  float q[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) q[i] = fminf(fmaxf(__fdiv_rn(f[i], s), -448.f), 448.f);
  *reinterpret_cast<uint2*>(xq + row * K + (int64_t)gcol * group_size + li * 8) =
      make_uint2(pack_fp8x4(q[0], q[1], q[2], q[3]), pack_fp8x4(q[4], q[5], q[6], q[7]));
}

// ---- fp8 per token -------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) per_token_quant_fp8_kernel(uint8_t* __restrict__ xq, float* __restrict__ scale,
                                                                  const T* __restrict__ x, int64_t K,
                                                                  int64_t x_row_stride) {
  __shared__ float sm[8];
  pdl_trigger();
  pdl_wait();
  const int64_t row = blockIdx.x;
  const T* xr = x + row * x_row_stride;
  float amax = 0.f;
  for (int64_t c = threadIdx.x * 8; c < K; c += 256 * 8) {
    float f[8];
    load8<T>(xr + c, f);
#pragma unroll
    for (int i = 0; i < 8; ++i) amax = fmaxf(amax, fabsf(f[i]));
  }
  amax = block_max<256>(amax, sm);
  const float s = __fdiv_rn(amax, 448.f);
  const float inv = (s == 0.f) ? 0.f : __fdiv_rn(1.f, s);
  if (threadIdx.x == 0) scale[row] = s;
  uint8_t* qr = xq + row * K;
  for (int64_t c = threadIdx.x * 8; c < K; c += 256 * 8) {
    float f[8], q[8];
    load8<T>(xr + c, f);
#pragma unroll
    for (int i = 0; i < 8; ++i) q[i] = fminf(fmaxf(__fmul_rn(f[i], inv), -448.f), 448.f);
    *reinterpret_cast<uint2*>(qr + c) =
        make_uint2(pack_fp8x4(q[0], q[1], q[2], q[3]), pack_fp8x4(q[4], q[5], q[6], q[7]));
  }
}

}  // namespace b200

using namespace b200;

B200_API int b200_per_token_quant_int8(int8_t* xq, float* scale, const void* x, int64_t M, int64_t K,
                                         int64_t x_row_stride, int dtype, void* stream) {
  B200_REQUIRE(M >= 0 && K > 0, "per_token_quant_int8: bad sizes");
  if (M == 0) return B200_OK;
  B200_REQUIRE(xq && scale && x, "per_token_quant_int8: null pointer");
  B200_REQUIRE(K % 8 == 0 && x_row_stride % 8 == 0, "per_token_quant_int8: K and row stride must be multiples of 8");
  B200_REQUIRE((uintptr_t)x % 16 == 0 && (uintptr_t)xq % 8 == 0, "per_token_quant_int8: alignment");
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == B200_BF16)
    return launch_k(per_token_quant_int8_kernel<__nv_bfloat16>, dim3((unsigned)M), dim3(256), 0, st, 1, true, xq, scale,
                    (const __nv_bfloat16*)x, K, x_row_stride);
  if (dtype == B200_F16)
    return launch_k(per_token_quant_int8_kernel<__half>, dim3((unsigned)M), dim3(256), 0, st, 1, true, xq, scale,
                    (const __half*)x, K, x_row_stride);
  return fail(B200_ERR_UNSUPPORTED, "per_token_quant_int8: dtype %d not supported", dtype);
}

B200_API int b200_per_token_group_quant_fp8(void* xq, float* scales, const void* x, int64_t M, int64_t K,
                                              int64_t x_row_stride, int group_size, float eps,
                                              int64_t scale_row_stride, int64_t scale_col_stride, int dtype,
                                              void* stream) {
  B200_REQUIRE(M >= 0 && K > 0, "per_token_group_quant_fp8: bad sizes");
  if (M == 0) return B200_OK;
  B200_REQUIRE(xq && scales && x, "per_token_group_quant_fp8: null pointer");
  B200_REQUIRE(group_size == 32 || group_size == 64 || group_size == 128 || group_size == 256,
               "per_token_group_quant_fp8: group_size %d not in {32,64,128,256}", group_size);
  B200_REQUIRE(K % group_size == 0, "per_token_group_quant_fp8: K %lld not divisible by group_size %d",
               (long long)K, group_size);
  B200_REQUIRE(x_row_stride % 8 == 0 && (uintptr_t)x % 16 == 0 && (uintptr_t)xq % 8 == 0,
               "per_token_group_quant_fp8: alignment");
  const int groups_per_warp = 32 / (group_size / 8);
  const int64_t total_groups = M * (K / group_size);
  const int64_t warps = ceil_div(total_groups, groups_per_warp);
  const unsigned blocks = (unsigned)ceil_div(warps, 8);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == B200_BF16)
    return launch_k(per_token_group_quant_fp8_kernel<__nv_bfloat16>, dim3(blocks), dim3(256), 0, st, 1, true, (uint8_t*)xq,
                    scales, (const __nv_bfloat16*)x, M, K, x_row_stride, group_size, eps, scale_row_stride, scale_col_stride);
  if (dtype == B200_F16)
    return launch_k(per_token_group_quant_fp8_kernel<__half>, dim3(blocks), dim3(256), 0, st, 1, true, (uint8_t*)xq, scales,
                    (const __half*)x, M, K, x_row_stride, group_size, eps, scale_row_stride, scale_col_stride);
  return fail(B200_ERR_UNSUPPORTED, "per_token_group_quant_fp8: dtype %d not supported", dtype);
}

B200_API int b200_per_token_quant_fp8(void* xq, float* scale, const void* x, int64_t M, int64_t K,
                                        int64_t x_row_stride, int dtype, void* stream) {
  B200_REQUIRE(M >= 0 && K > 0, "per_token_quant_fp8: bad sizes");
  if (M == 0) return B200_OK;
  B200_REQUIRE(xq && scale && x, "per_token_quant_fp8: null pointer");
  B200_REQUIRE(K % 8 == 0 && x_row_stride % 8 == 0 && (uintptr_t)x % 16 == 0 && (uintptr_t)xq % 8 == 0,
               "per_token_quant_fp8: alignment");
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == B200_BF16)
    return launch_k(per_token_quant_fp8_kernel<__nv_bfloat16>, dim3((unsigned)M), dim3(256), 0, st, 1, true, (uint8_t*)xq, scale,
                    (const __nv_bfloat16*)x, K, x_row_stride);
  if (dtype == B200_F16)
    return launch_k(per_token_quant_fp8_kernel<__half>, dim3((unsigned)M), dim3(256), 0, st, 1, true, (uint8_t*)xq, scale,
                    (const __half*)x, K, x_row_stride);
  return fail(B200_ERR_UNSUPPORTED, "per_token_quant_fp8: dtype %d not supported", dtype);
}
