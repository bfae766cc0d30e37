// Shared pieces of the tcgen05 GEMM kernels (gemm_umma.cu, gemm_fp8_block.cu).
#pragma once

#include "sm100.cuh"

namespace b200 {

using namespace sm100;

enum GemmMode { kModeI8 = 0, kModeF8Chan = 1, kModeF8Block = 2, kModeBF16 = 3, kModeF16 = 4 };  // 3/4: un-scaled 16-bit operands

struct GemmArgs {
  void* out;             // [M, N] bf16/f16
  const float* scale_m;  // per token: [M] (I8 / F8Chan) or [M, K/128] with strides (F8Block)
  const float* scale_n;  // per channel: [N] or [N/128, K/128] with strides (F8Block)
  const void* bias;      // [N] out dtype or null
  int32_t* acc_out;      // [M, N] raw int32 accumulators (I8 only) or null
  int M, N;
  int num_kb;        // K blocks of 128 bytes (ceil)
  int kb_per_split;  // K blocks per cluster rank
  int splits;        // cluster size along z (1, 2, 4, 8)
  int out_is_bf16;
  int m_blocks;     // stream-K: token blocks of BN per 128-channel weight tile (>= 1)
  int64_t sm_rs, sm_cs;  // scale_m[m * sm_rs + kb * sm_cs]          (F8Block)
  int64_t sn_rs, sn_cs;  // scale_n[(n / 128) * sn_rs + kb * sn_cs]   (F8Block)
  unsigned long long* trace;  // diagnostics (b200_debug_set_trace) or null
  int trace_slots;
  int debug;  // B200_GEMM_DEBUG bit mask (timing experiments only: results are wrong when set)
};

__device__ __forceinline__ unsigned long long* trace_slot(const GemmArgs& g) {
  const int cta = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
  return (g.trace != nullptr && cta < g.trace_slots) ? g.trace + (size_t)cta * 16 : nullptr;
}

constexpr int kGemmThreads = 256;
constexpr int kBK = 128;  // bytes of K per pipeline stage = one 128B swizzle row

__device__ __forceinline__ float load_bias(const void* bias, int n, int is_bf16) {
  return is_bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(bias)[n])
                 : __half2float(reinterpret_cast<const __half*>(bias)[n]);
}

__device__ __forceinline__ void store_out(void* out, int64_t idx, float v, int is_bf16) {
  if (is_bf16) reinterpret_cast<__nv_bfloat16*>(out)[idx] = __float2bfloat16_rn(v);
  else reinterpret_cast<__half*>(out)[idx] = __float2half_rn(v);
}

constexpr int kSkFlagBytes = 1024;  // up to 256 CTAs

__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 128;\n" ::: "memory"); }


// gemm_fp8_block.cu
int launch_block_fp8_streamk(GemmArgs g, const void* a, const void* b, int64_t M, int64_t N, int64_t K, void* ws,
                             cudaStream_t st);
int launch_block_fp8_prefill(GemmArgs g, const void* a, const void* b, int64_t M, int64_t N, int64_t K, cudaStream_t st);

}  // namespace b200
