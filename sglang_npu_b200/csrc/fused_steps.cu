// Fused element-wise steps either side of the quantised linears and the KV write:
//   b200_rmsnorm_quant_int8      (fused_add_)rmsnorm  -> per-token int8 quant   (feeds qkv_proj / gate_up_proj)
//   b200_silu_and_mul_quant_int8 silu_and_mul         -> per-token int8 quant   (feeds down_proj)
//   b200_rope_neox_set_kv        NeoX RoPE on q,k (in place) + set_kv_buffer of (rotated k, v)
//   b200_rmsnorm_group_quant_fp8 / b200_silu_and_mul_group_quant_fp8: the same two pairs for the block-scaled FP8
//                                linears (per-(token, 128) e4m3 quant: the absmax is local to 16 lanes)
// Each one is bit-identical to running the two un-fused kernels back to back (elementwise.cu then quant.cu /
// kv_write.cu): the intermediate activation is rounded to the activation dtype before the absmax and the
// int8 rounding, exactly as when it round-trips through HBM.  What the fusion removes is a launch and an
// activation round-trip per linear — at decode sizes these steps are latency-bound, not bandwidth-bound.
// Reference kernels they stand in for: sgl-kernel/csrc/elementwise/{fused_add_rms_norm_kernel,activation,
// rope}.cu, python/sglang/srt/layers/quantization/int8_kernel.py:20-83, mem_cache/memory_pool.py:369-407.
#include <cuda_fp8.h>

#include "common.cuh"

namespace b200 {

namespace {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_maxf(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

template <typename T>
__device__ __forceinline__ void load8f(const T* p, float (&f)[8]) {
  const int4 v = *reinterpret_cast<const int4*>(p);
  const uint32_t* u = reinterpret_cast<const uint32_t*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 x = unpack2<T>(u[i]);
    f[2 * i] = x.x;
    f[2 * i + 1] = x.y;
  }
}

// int8 codes of 8 values (same rule as per_token_quant_int8_kernel: rn(x * (rcp(absmax) * 127)))
__device__ __forceinline__ uint2 quant8(const float (&f)[8], float inv) {
  uint32_t lo = 0, hi = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int a = __float2int_rn(__fmul_rn(f[i], inv));
    const int b = __float2int_rn(__fmul_rn(f[4 + i], inv));
    lo |= (uint32_t)(a & 0xff) << (8 * i);
    hi |= (uint32_t)(b & 0xff) << (8 * i);
  }
  return make_uint2(lo, hi);
}

}  // namespace

// ---- (fused add +) RMSNorm -> int8 -------------------------------------------------------------------------
// One CTA per token row.  y = T(residual' * rstd * w) is kept in shared memory as T-rounded floats; y_out
// (optional) receives it for callers that also need the normalised activation.
template <typename T>
__global__ void __launch_bounds__(256) rmsnorm_quant_int8_kernel(int8_t* __restrict__ xq, float* __restrict__ scale,
                                                                 T* __restrict__ y_out, T* __restrict__ residual,
                                                                 const T* __restrict__ x, const T* __restrict__ w,
                                                                 int64_t H, float eps) {
  extern __shared__ float row_f[];  // H floats
  __shared__ float red[8];
  __shared__ float red2[8];
  pdl_trigger();
  pdl_wait();
  const int64_t row = blockIdx.x;
  const T* xr = x + row * H;
  T* rr = residual ? residual + row * H : nullptr;
  float ss = 0.f;
  for (int64_t c = threadIdx.x * 8; c < H; c += 256 * 8) {
    float f[8];
    load8f<T>(xr + c, f);
    if (rr) {
      float r[8];
      load8f<T>(rr + c, r);
      int4 nv;
      uint32_t* nu = reinterpret_cast<uint32_t*>(&nv);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        f[2 * i] += r[2 * i];
        f[2 * i + 1] += r[2 * i + 1];
        nu[i] = pack2<T>(f[2 * i], f[2 * i + 1]);
      }
      *reinterpret_cast<int4*>(rr + c) = nv;  // residual = T(x + residual)
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      row_f[c + i] = f[i];
      ss += f[i] * f[i];
    }
  }
  ss = warp_sum(ss);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  float tot = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) tot += red[i];
  const float rstd = rsqrtf(tot / (float)H + eps);
  float amax = 0.f;
  for (int64_t c = threadIdx.x * 8; c < H; c += 256 * 8) {
    float wv[8];
    load8f<T>(w + c, wv);
    int4 ov;
    uint32_t* ou = reinterpret_cast<uint32_t*>(&ov);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      ou[i] = pack2<T>(row_f[c + 2 * i] * rstd * wv[2 * i], row_f[c + 2 * i + 1] * rstd * wv[2 * i + 1]);
      const float2 y = unpack2<T>(ou[i]);  // the T-rounded activation the un-fused quant kernel would read
      row_f[c + 2 * i] = y.x;
      row_f[c + 2 * i + 1] = y.y;
      amax = fmaxf(amax, fmaxf(fabsf(y.x), fabsf(y.y)));
    }
    if (y_out) *reinterpret_cast<int4*>(y_out + row * H + c) = ov;
  }
  amax = warp_maxf(amax);
  if ((threadIdx.x & 31) == 0) red2[threadIdx.x >> 5] = amax;
  __syncthreads();
  amax = red2[0];
#pragma unroll
  for (int i = 1; i < 8; ++i) amax = fmaxf(amax, red2[i]);
  amax = fmaxf(amax, 1e-10f);
  const float inv = __fmul_rn(__frcp_rn(amax), 127.f);
  if (threadIdx.x == 0) scale[row] = __fdiv_rn(amax, 127.f);
  int8_t* qr = xq + row * H;
  for (int64_t c = threadIdx.x * 8; c < H; c += 256 * 8) {
    float f[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) f[i] = row_f[c + i];
    *reinterpret_cast<uint2*>(qr + c) = quant8(f, inv);
  }
}

// ---- silu_and_mul -> int8 ------------------------------------------------------------------------------------
// A row of 2H gate|up values is handled by a cluster of CL CTAs (each H/CL columns) so that a 64-token decode
// batch still spreads over 128+ SMs; the row absmax is exchanged through distributed shared memory.
template <typename T>
__global__ void __launch_bounds__(256) silu_mul_quant_int8_kernel(int8_t* __restrict__ xq, float* __restrict__ scale,
                                                                  const T* __restrict__ x, int64_t H, int cols_per_cta,
                                                                  int cluster) {
  extern __shared__ uint32_t act[];  // cols_per_cta / 2 packed T pairs
  __shared__ float red[8];
  __shared__ float peer_max[8];
  pdl_trigger();
  pdl_wait();
  const int64_t row = blockIdx.x;
  const int rank = blockIdx.z;
  const int64_t c_lo = (int64_t)rank * cols_per_cta;
  const int64_t c_hi = min((int64_t)H, c_lo + cols_per_cta);
  const T* g = x + row * 2 * H;
  const T* u = g + H;
  float amax = 0.f;
  for (int64_t c = c_lo + threadIdx.x * 8; c < c_hi; c += 256 * 8) {
    float gv[8], uv[8];
    load8f<T>(g + c, gv);
    load8f<T>(u + c, uv);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      // F.silu(gate) is rounded to T before the multiply (activation.py:60-62 forward_native)
      const float s0 = DT<T>::to_f(DT<T>::from_f(gv[2 * i] / (1.f + __expf(-gv[2 * i]))));
      const float s1 = DT<T>::to_f(DT<T>::from_f(gv[2 * i + 1] / (1.f + __expf(-gv[2 * i + 1]))));
      const uint32_t p = pack2<T>(s0 * uv[2 * i], s1 * uv[2 * i + 1]);
      act[((c - c_lo) >> 1) + i] = p;
      const float2 y = unpack2<T>(p);
      amax = fmaxf(amax, fmaxf(fabsf(y.x), fabsf(y.y)));
    }
  }
  amax = warp_maxf(amax);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = amax;
  __syncthreads();
  amax = red[0];
#pragma unroll
  for (int i = 1; i < 8; ++i) amax = fmaxf(amax, red[i]);
  if (cluster > 1) {
    if (threadIdx.x < cluster) {
      uint32_t remote;
      const uint32_t local = smem_u32(&peer_max[rank]);
      asm volatile("mapa.shared::cluster.u32 %0, %1, %2;\n" : "=r"(remote) : "r"(local), "r"((uint32_t)threadIdx.x));
      asm volatile("st.shared::cluster.f32 [%0], %1;\n" ::"r"(remote), "f"(amax) : "memory");
    }
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;\n" ::: "memory");
    amax = peer_max[0];
    for (int i = 1; i < cluster; ++i) amax = fmaxf(amax, peer_max[i]);
  }
  amax = fmaxf(amax, 1e-10f);
  const float inv = __fmul_rn(__frcp_rn(amax), 127.f);
  if (threadIdx.x == 0 && rank == 0) scale[row] = __fdiv_rn(amax, 127.f);
  int8_t* qr = xq + row * H;
  for (int64_t c = c_lo + threadIdx.x * 8; c < c_hi; c += 256 * 8) {
    float f[8];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 y = unpack2<T>(act[((c - c_lo) >> 1) + i]);
      f[2 * i] = y.x;
      f[2 * i + 1] = y.y;
    }
    *reinterpret_cast<uint2*>(qr + c) = quant8(f, inv);
  }
}

// ---- NeoX RoPE (in place) + KV-cache write ---------------------------------------------------------------------
// One CTA per token.  Work items of 8 pairs: item -> (head, j0): x1 = head[j0..j0+7], x2 = head[half+j0..].
// Rotated k rows go both back to k (the reference rotates in place) and to k_cache[loc]; v rows are copied to
// v_cache[loc].  16-byte accesses throughout.
template <typename T>
__global__ void __launch_bounds__(256) rope_set_kv_kernel(T* __restrict__ q, T* __restrict__ k, const T* __restrict__ v,
                                                          const float* __restrict__ cos_sin, const int64_t* __restrict__ pos,
                                                          T* __restrict__ k_cache, T* __restrict__ v_cache,
                                                          const void* __restrict__ loc, int loc_is_int64, int num_q_heads,
                                                          int num_kv_heads, int head_dim, int rot_dim, int64_t q_stride,
                                                          int64_t k_stride, int64_t v_stride, int64_t kc_stride,
                                                          int64_t vc_stride) {
  pdl_trigger();
  pdl_wait();
  const int64_t tok = blockIdx.x;
  const int half = rot_dim / 2;
  const int items_per_head = half / 8;
  const float* cs = cos_sin + pos[tok] * rot_dim;
  const int64_t slot = k_cache ? load_index(loc, tok, loc_is_int64) : 0;
  const int total = (num_q_heads + num_kv_heads) * items_per_head;
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const int h = i / items_per_head;
    const int j0 = (i - h * items_per_head) * 8;
    const bool is_k = h >= num_q_heads;
    T* base = is_k ? k + tok * k_stride + (int64_t)(h - num_q_heads) * head_dim : q + tok * q_stride + (int64_t)h * head_dim;
    float x1[8], x2[8];
    load8f<T>(base + j0, x1);
    load8f<T>(base + half + j0, x2);
    const float4 c0 = *reinterpret_cast<const float4*>(cs + j0), c1 = *reinterpret_cast<const float4*>(cs + j0 + 4);
    const float4 s0 = *reinterpret_cast<const float4*>(cs + half + j0), s1 = *reinterpret_cast<const float4*>(cs + half + j0 + 4);
    const float c[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
    const float s[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
    int4 o1, o2;
    uint32_t* u1 = reinterpret_cast<uint32_t*>(&o1);
    uint32_t* u2 = reinterpret_cast<uint32_t*>(&o2);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      // same expression order as rope_neox_kernel (x1*c - x2*s, x2*c + x1*s)
      u1[e] = pack2<T>(x1[2 * e] * c[2 * e] - x2[2 * e] * s[2 * e], x1[2 * e + 1] * c[2 * e + 1] - x2[2 * e + 1] * s[2 * e + 1]);
      u2[e] = pack2<T>(x2[2 * e] * c[2 * e] + x1[2 * e] * s[2 * e], x2[2 * e + 1] * c[2 * e + 1] + x1[2 * e + 1] * s[2 * e + 1]);
    }
    *reinterpret_cast<int4*>(base + j0) = o1;
    *reinterpret_cast<int4*>(base + half + j0) = o2;
    if (is_k && k_cache) {
      T* dst = k_cache + slot * kc_stride + (int64_t)(h - num_q_heads) * head_dim;
      *reinterpret_cast<int4*>(dst + j0) = o1;
      *reinterpret_cast<int4*>(dst + half + j0) = o2;
    }
  }
  if (k_cache) {
    // non-rotated tail of k (rot_dim < head_dim) and all of v
    const int tail_chunks = (head_dim - rot_dim) / 8;
    for (int i = threadIdx.x; i < num_kv_heads * tail_chunks; i += blockDim.x) {
      const int h = i / tail_chunks;
      const int c = rot_dim + (i - h * tail_chunks) * 8;
      *reinterpret_cast<int4*>(k_cache + slot * kc_stride + (int64_t)h * head_dim + c) =
          *reinterpret_cast<const int4*>(k + tok * k_stride + (int64_t)h * head_dim + c);
    }
    const int v_chunks = num_kv_heads * head_dim / 8;
    for (int i = threadIdx.x; i < v_chunks; i += blockDim.x)
      *reinterpret_cast<int4*>(v_cache + slot * vc_stride + (int64_t)i * 8) =
          *reinterpret_cast<const int4*>(v + tok * v_stride + (int64_t)i * 8);
  }
}


// ---- block-FP8 activation quantisation fused behind the producer ---------------------------------------------------
// Same arithmetic as per_token_group_quant_fp8_kernel (quant.cu): s = max(absmax, eps) / 448, q = sat_e4m3(y / s), on
// the T-rounded activation the un-fused kernel would have read back from HBM; a 128-column group = 16 lanes x 8.
__device__ __forceinline__ uint32_t fs_pack_fp8x4(float a, float b, float c, float d) {
  const __nv_fp8x2_storage_t lo = __nv_cvt_float2_to_fp8x2(make_float2(a, b), __NV_SATFINITE, __NV_E4M3);
  const __nv_fp8x2_storage_t hi = __nv_cvt_float2_to_fp8x2(make_float2(c, d), __NV_SATFINITE, __NV_E4M3);
  return (uint32_t)lo | ((uint32_t)hi << 16);
}
__device__ __forceinline__ uint2 group_quant8(const float (&y)[8], float eps, float* scale_out, bool write_scale) {
  float amax = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) amax = fmaxf(amax, fabsf(y[i]));
  const uint32_t half_mask = 0xFFFFu << (threadIdx.x & 16);  // the 16 lanes of this group (the other half-warp may be idle)
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(half_mask, amax, o));
  amax = fmaxf(amax, eps);
  const float s = __fdiv_rn(amax, 448.f);
  if (write_scale) *scale_out = s;
  float q[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) q[i] = fminf(fmaxf(__fdiv_rn(y[i], s), -448.f), 448.f);
  return make_uint2(fs_pack_fp8x4(q[0], q[1], q[2], q[3]), fs_pack_fp8x4(q[4], q[5], q[6], q[7]));
}

// One CTA per token row, H % 128 == 0.  Thread t owns columns [8t, 8t+8) (+2048 per pass): a group is 16 consecutive
// lanes of one warp, so the group absmax is four shuffles.
template <typename T>
__global__ void __launch_bounds__(256) rmsnorm_group_quant_fp8_kernel(uint8_t* __restrict__ xq, float* __restrict__ scales,
                                                                      T* __restrict__ residual, const T* __restrict__ x,
                                                                      const T* __restrict__ w, int64_t H, float eps,
                                                                      float quant_eps) {
  extern __shared__ float row_f[];  // H floats
  __shared__ float red[8];
  pdl_trigger();
  pdl_wait();
  const int64_t row = blockIdx.x;
  const T* xr = x + row * H;
  T* rr = residual ? residual + row * H : nullptr;
  float ss = 0.f;
  for (int64_t c = threadIdx.x * 8; c < H; c += 256 * 8) {
    float f[8];
    load8f<T>(xr + c, f);
    if (rr) {
      float r[8];
      load8f<T>(rr + c, r);
      int4 nv;
      uint32_t* nu = reinterpret_cast<uint32_t*>(&nv);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        f[2 * i] += r[2 * i];
        f[2 * i + 1] += r[2 * i + 1];
        nu[i] = pack2<T>(f[2 * i], f[2 * i + 1]);
      }
      *reinterpret_cast<int4*>(rr + c) = nv;  // residual = T(x + residual)
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      row_f[c + i] = f[i];
      ss += f[i] * f[i];
    }
  }
  ss = warp_sum(ss);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  float tot = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) tot += red[i];
  const float rstd = rsqrtf(tot / (float)H + eps);
  const int64_t groups = H / 128;
  for (int64_t c = threadIdx.x * 8; c < H; c += 256 * 8) {  // H % 128 == 0: a half-warp is all in or all out
    float wv[8], y[8];
    load8f<T>(w + c, wv);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 t2 = unpack2<T>(pack2<T>(row_f[c + 2 * i] * rstd * wv[2 * i], row_f[c + 2 * i + 1] * rstd * wv[2 * i + 1]));
      y[2 * i] = t2.x;
      y[2 * i + 1] = t2.y;
    }
    const int64_t g = c >> 7;
    *reinterpret_cast<uint2*>(xq + row * H + c) = group_quant8(y, quant_eps, scales + row * groups + g, (threadIdx.x & 15) == 0);
  }
}

// x [M, 2H] gate | up -> e4m3 [M, H] + scales [M, H / 128].  A warp quantises two groups; grid-stride over groups.
template <typename T>
__global__ void __launch_bounds__(256) silu_mul_group_quant_fp8_kernel(uint8_t* __restrict__ xq, float* __restrict__ scales,
                                                                       const T* __restrict__ x, int64_t M, int64_t H,
                                                                       float quant_eps) {
  pdl_trigger();
  pdl_wait();
  const int64_t groups = H / 128;
  const int64_t total = M * groups;  // groups overall; half-warp hw handles group hw, hw + stride, ...
  const int64_t hw0 = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 4;
  const int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 4;
  const int li = threadIdx.x & 15;
  // both half-warps of a warp must run the same number of iterations (full-mask shuffles): pad with an inactive pass
  const int64_t iters = (total + stride - 1) / stride;
  for (int64_t it = 0; it < iters; ++it) {
    const int64_t gi = hw0 + it * stride;
    const bool active = gi < total;
    const int64_t row = active ? gi / groups : 0;
    const int64_t g = active ? gi - row * groups : 0;
    const int64_t c = g * 128 + li * 8;
    float y[8];
    if (active) {
      const int4 gv = *reinterpret_cast<const int4*>(x + row * 2 * H + c);
      const int4 uv = *reinterpret_cast<const int4*>(x + row * 2 * H + H + c);
      const uint32_t* gu = reinterpret_cast<const uint32_t*>(&gv);
      const uint32_t* uu = reinterpret_cast<const uint32_t*>(&uv);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float2 g2 = unpack2<T>(gu[k]);
        const float2 u2 = unpack2<T>(uu[k]);
        // same arithmetic as silu_and_mul_kernel (elementwise.cu): F.silu(gate) rounded to T before the multiply
        const float s0 = DT<T>::to_f(DT<T>::from_f(g2.x / (1.f + __expf(-g2.x))));
        const float s1 = DT<T>::to_f(DT<T>::from_f(g2.y / (1.f + __expf(-g2.y))));
        const float2 t2 = unpack2<T>(pack2<T>(s0 * u2.x, s1 * u2.y));  // the T-rounded activation
        y[2 * k] = t2.x;
        y[2 * k + 1] = t2.y;
      }
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k) y[k] = 0.f;
    }
    float sdummy;
    const uint2 q = group_quant8(y, quant_eps, active ? scales + row * groups + g : &sdummy, li == 0);
    if (active) *reinterpret_cast<uint2*>(xq + row * H + c) = q;
  }
}

}  // namespace b200

using namespace b200;

B200_API int b200_rmsnorm_quant_int8(int8_t* xq, float* scale, void* y_out, void* residual, const void* x,
                                       const void* weight, int64_t M, int64_t H, float eps, int dtype, void* stream) {
  B200_REQUIRE(M >= 0 && H > 0, "rmsnorm_quant_int8: bad sizes");
  if (M == 0) return B200_OK;
  B200_REQUIRE(xq && scale && x && weight, "rmsnorm_quant_int8: null pointer");
  B200_REQUIRE(H % 8 == 0 && H <= 16384, "rmsnorm_quant_int8: H %lld must be a multiple of 8 and <= 16384", (long long)H);
  B200_REQUIRE(((uintptr_t)x | (uintptr_t)weight | (uintptr_t)residual | (uintptr_t)y_out) % 16 == 0 && (uintptr_t)xq % 8 == 0,
               "rmsnorm_quant_int8: alignment");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t smem = (size_t)H * sizeof(float);
  if (dtype == B200_BF16) {
    auto kern = rmsnorm_quant_int8_kernel<__nv_bfloat16>;
    if (smem > 48 * 1024) B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    return launch_k(kern, dim3((unsigned)M), dim3(256), smem, st, 1, true, xq, scale, (__nv_bfloat16*)y_out,
                    (__nv_bfloat16*)residual, (const __nv_bfloat16*)x, (const __nv_bfloat16*)weight, H, eps);
  }
  if (dtype == B200_F16) {
    auto kern = rmsnorm_quant_int8_kernel<__half>;
    if (smem > 48 * 1024) B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    return launch_k(kern, dim3((unsigned)M), dim3(256), smem, st, 1, true, xq, scale, (__half*)y_out, (__half*)residual,
                    (const __half*)x, (const __half*)weight, H, eps);
  }
  return fail(B200_ERR_UNSUPPORTED, "rmsnorm_quant_int8: dtype %d not supported", dtype);
}

B200_API int b200_silu_and_mul_quant_int8(int8_t* xq, float* scale, const void* x, int64_t M, int64_t H, int dtype,
                                            void* stream) {
  B200_REQUIRE(M >= 0 && H > 0, "silu_and_mul_quant_int8: bad sizes");
  if (M == 0) return B200_OK;
  B200_REQUIRE(xq && scale && x, "silu_and_mul_quant_int8: null pointer");
  B200_REQUIRE(H % 8 == 0, "silu_and_mul_quant_int8: H must be a multiple of 8");
  B200_REQUIRE((uintptr_t)x % 16 == 0 && (uintptr_t)xq % 8 == 0, "silu_and_mul_quant_int8: alignment");
  cudaStream_t st = (cudaStream_t)stream;
  // spread a row over a cluster while the batch alone cannot fill the SMs; each CTA keeps <= 16K columns in smem
  int cluster = 1;
  while (cluster < 8 && (M * cluster < sm_count() || ceil_div(H, cluster) > 16384) && ceil_div(H, cluster * 2) >= 2048) cluster *= 2;
  int cols = (int)ceil_div(ceil_div(H, cluster), 8) * 8;
  B200_REQUIRE(cols <= 32768, "silu_and_mul_quant_int8: H %lld too large", (long long)H);
  const size_t smem = (size_t)cols * 2;
  dim3 grid((unsigned)M, 1, (unsigned)cluster);
  if (dtype == B200_BF16) {
    auto kern = silu_mul_quant_int8_kernel<__nv_bfloat16>;
    if (smem > 48 * 1024) B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    return launch_k(kern, grid, dim3(256), smem, st, cluster, true, xq, scale, (const __nv_bfloat16*)x, H, cols, cluster);
  }
  if (dtype == B200_F16) {
    auto kern = silu_mul_quant_int8_kernel<__half>;
    if (smem > 48 * 1024) B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    return launch_k(kern, grid, dim3(256), smem, st, cluster, true, xq, scale, (const __half*)x, H, cols, cluster);
  }
  return fail(B200_ERR_UNSUPPORTED, "silu_and_mul_quant_int8: dtype %d not supported", dtype);
}

B200_API int b200_rope_neox_set_kv(void* q, void* k, const void* v, const float* cos_sin_cache, const int64_t* positions,
                                     void* k_cache, void* v_cache, const void* loc, int loc_is_int64, int64_t num_tokens,
                                     int num_q_heads, int num_kv_heads, int head_dim, int rot_dim, int64_t q_token_stride,
                                     int64_t k_token_stride, int64_t v_token_stride, int64_t k_cache_token_stride,
                                     int64_t v_cache_token_stride, int dtype, void* stream) {
  B200_REQUIRE(num_tokens >= 0, "rope_neox_set_kv: bad sizes");
  if (num_tokens == 0) return B200_OK;
  B200_REQUIRE(q && k && cos_sin_cache && positions, "rope_neox_set_kv: null pointer");
  B200_REQUIRE((k_cache == nullptr) == (v_cache == nullptr), "rope_neox_set_kv: k_cache and v_cache go together");
  B200_REQUIRE(k_cache == nullptr || (v && loc), "rope_neox_set_kv: v / loc missing");
  B200_REQUIRE(rot_dim % 16 == 0 && rot_dim <= head_dim && head_dim % 8 == 0,
               "rope_neox_set_kv: rot_dim %d must be a multiple of 16 and <= head_dim %d", rot_dim, head_dim);
  B200_REQUIRE((q_token_stride | k_token_stride | v_token_stride | k_cache_token_stride | v_cache_token_stride) % 8 == 0 &&
                   ((uintptr_t)q | (uintptr_t)k | (uintptr_t)v | (uintptr_t)k_cache | (uintptr_t)v_cache |
                    (uintptr_t)cos_sin_cache) % 16 == 0,
               "rope_neox_set_kv: 16-byte alignment of rows");
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == B200_BF16)
    return launch_k(rope_set_kv_kernel<__nv_bfloat16>, dim3((unsigned)num_tokens), dim3(256), 0, st, 1, true,
                    (__nv_bfloat16*)q, (__nv_bfloat16*)k, (const __nv_bfloat16*)v, cos_sin_cache, positions,
                    (__nv_bfloat16*)k_cache, (__nv_bfloat16*)v_cache, loc, loc_is_int64, num_q_heads, num_kv_heads, head_dim,
                    rot_dim, q_token_stride, k_token_stride, v_token_stride, k_cache_token_stride, v_cache_token_stride);
  if (dtype == B200_F16)
    return launch_k(rope_set_kv_kernel<__half>, dim3((unsigned)num_tokens), dim3(256), 0, st, 1, true, (__half*)q, (__half*)k,
                    (const __half*)v, cos_sin_cache, positions, (__half*)k_cache, (__half*)v_cache, loc, loc_is_int64,
                    num_q_heads, num_kv_heads, head_dim, rot_dim, q_token_stride, k_token_stride, v_token_stride,
                    k_cache_token_stride, v_cache_token_stride);
  return fail(B200_ERR_UNSUPPORTED, "rope_neox_set_kv: dtype %d not supported", dtype);
}

B200_API int b200_rmsnorm_group_quant_fp8(void* xq, float* scales, void* residual, const void* x, const void* weight,
                                            int64_t M, int64_t H, float eps, float quant_eps, int dtype, void* stream) {
  B200_REQUIRE(M >= 0 && H > 0, "rmsnorm_group_quant_fp8: bad sizes");
  if (M == 0) return B200_OK;
  B200_REQUIRE(xq && scales && x && weight, "rmsnorm_group_quant_fp8: null pointer");
  B200_REQUIRE(H % 128 == 0 && H <= 16384, "rmsnorm_group_quant_fp8: H %lld must be a multiple of 128 and <= 16384",
               (long long)H);
  B200_REQUIRE(((uintptr_t)x | (uintptr_t)weight | (uintptr_t)residual) % 16 == 0 && (uintptr_t)xq % 8 == 0,
               "rmsnorm_group_quant_fp8: alignment");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t smem = (size_t)H * sizeof(float);
  if (dtype == B200_BF16) {
    auto kern = rmsnorm_group_quant_fp8_kernel<__nv_bfloat16>;
    if (smem > 48 * 1024) B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    return launch_k(kern, dim3((unsigned)M), dim3(256), smem, st, 1, true, (uint8_t*)xq, scales, (__nv_bfloat16*)residual,
                    (const __nv_bfloat16*)x, (const __nv_bfloat16*)weight, H, eps, quant_eps);
  }
  if (dtype == B200_F16) {
    auto kern = rmsnorm_group_quant_fp8_kernel<__half>;
    if (smem > 48 * 1024) B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    return launch_k(kern, dim3((unsigned)M), dim3(256), smem, st, 1, true, (uint8_t*)xq, scales, (__half*)residual,
                    (const __half*)x, (const __half*)weight, H, eps, quant_eps);
  }
  return fail(B200_ERR_UNSUPPORTED, "rmsnorm_group_quant_fp8: dtype %d not supported", dtype);
}

B200_API int b200_silu_and_mul_group_quant_fp8(void* xq, float* scales, const void* x, int64_t M, int64_t H,
                                                 float quant_eps, int dtype, void* stream) {
  B200_REQUIRE(M >= 0 && H > 0, "silu_and_mul_group_quant_fp8: bad sizes");
  if (M == 0) return B200_OK;
  B200_REQUIRE(xq && scales && x, "silu_and_mul_group_quant_fp8: null pointer");
  B200_REQUIRE(H % 128 == 0, "silu_and_mul_group_quant_fp8: H must be a multiple of 128");
  B200_REQUIRE((uintptr_t)x % 16 == 0 && (uintptr_t)xq % 8 == 0, "silu_and_mul_group_quant_fp8: alignment");
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t half_warps = M * (H / 128);
  int64_t blocks = ceil_div(half_warps, 16);
  const int64_t cap = (int64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  if (dtype == B200_BF16)
    return launch_k(silu_mul_group_quant_fp8_kernel<__nv_bfloat16>, dim3((unsigned)blocks), dim3(256), 0, st, 1, true,
                    (uint8_t*)xq, scales, (const __nv_bfloat16*)x, M, H, quant_eps);
  if (dtype == B200_F16)
    return launch_k(silu_mul_group_quant_fp8_kernel<__half>, dim3((unsigned)blocks), dim3(256), 0, st, 1, true, (uint8_t*)xq,
                    scales, (const __half*)x, M, H, quant_eps);
  return fail(B200_ERR_UNSUPPORTED, "silu_and_mul_group_quant_fp8: dtype %d not supported", dtype);
}
