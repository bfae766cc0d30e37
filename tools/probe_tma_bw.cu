// Bandwidth probe: how fast can 148 SMs pull a weight matrix out of HBM with different TMA access shapes?
// (standalone diagnostic, not part of the library:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o
//  gpurun_out/probe_tma_bw tools/probe_tma_bw.cu -lcuda ; ./gpurun_out/probe_tma_bw)
//   mode 0: 2-D tiled box [rows x 128 B] out of a row-major byte matrix with pitch P (the int8 GEMM's weight tile)
//   mode 1: 1-D bulk copy of one contiguous chunk (what a pre-tiled weight layout would allow)
//   mode 2: 2-D box [64 rows x 64 B], pitch P (the AWQ packed-word tile)
// Every CTA owns a strip and streams it through a STAGES-deep ring; nothing is computed.
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                              \
  do {                                                                                     \
    cudaError_t e = (x);                                                                   \
    if (e != cudaSuccess) {                                                                \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__);       \
      exit(1);                                                                             \
    }                                                                                      \
  } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t c) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(c));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok = 0;
  while (!ok) {
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                 : "=r"(ok)
                 : "r"(bar), "r"(parity)
                 : "memory");
  }
}
__device__ __forceinline__ void tma_2d(uint32_t dst, const CUtensorMap* m, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
      "l"((uint64_t)m), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void bulk_1d(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}

// strip = blockIdx.x: mode 0/2: rows [strip*box_rows, +box_rows) (mode 0) or byte columns [strip*64, +64) (mode 2);
// steps = number of boxes along the streaming dimension
template <int STAGES>
__global__ void __launch_bounds__(64) probe(const __grid_constant__ CUtensorMap tmap, const uint8_t* base, int mode,
                                            int steps, int box_bytes, int box_rows, int strips) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bars[STAGES];
  const uint32_t sbase = (smem_u32(smem) + 1023u) & ~1023u;
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) mbar_init(smem_u32(&bars[s]), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  for (int strip = blockIdx.x; strip < strips; strip += gridDim.x) {
    // the ring is drained at the end of each strip, so parities restart consistently
    for (int i = 0; i < steps + STAGES - 1; ++i) {
      if (i < steps) {
        const int s = i % STAGES;
        const uint32_t bar = smem_u32(&bars[s]);
        mbar_expect_tx(bar, box_bytes);
        const uint32_t dst = sbase + s * box_bytes;
        if (mode == 0) tma_2d(dst, &tmap, bar, i * 128, strip * box_rows);
        else if (mode == 2) tma_2d(dst, &tmap, bar, strip * 64, i * 64);
        else bulk_1d(dst, base + ((size_t)strip * steps + i) * box_bytes, box_bytes, bar);
      }
      const int j = i - (STAGES - 1);
      if (j >= 0) {
        const int s = j % STAGES;
        mbar_wait(smem_u32(&bars[s]), (uint32_t)((j / STAGES) & 1));
      }
    }
    // re-init barriers between strips (single thread, nothing in flight)
    for (int s = 0; s < STAGES; ++s) {
      asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(smem_u32(&bars[s])) : "memory");
      mbar_init(smem_u32(&bars[s]), 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
}

typedef CUresult (*encode_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static CUtensorMap make_map(encode_t enc, void* base, uint64_t rows, uint64_t row_bytes, uint32_t box_b, uint32_t box_r,
                            CUtensorMapSwizzle sw) {
  CUtensorMap m;
  cuuint64_t dims[2] = {row_bytes, rows};
  cuuint64_t strides[1] = {row_bytes};
  cuuint32_t box[2] = {box_b, box_r};
  cuuint32_t es[2] = {1, 1};
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    printf("encode failed %d\n", (int)r);
    exit(1);
  }
  return m;
}

template <int STAGES>
static void run(const char* name, const CUtensorMap& map, const uint8_t* base, int mode, int steps, int box_bytes,
                int box_rows, int strips, int grid, size_t total_bytes, uint8_t* flush, size_t flush_bytes,
                bool one_cta_per_sm = false) {
  const int smem = one_cta_per_sm ? 120 * 1024 : STAGES * box_bytes + 1024;
  CK(cudaFuncSetAttribute(probe<STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  float best = 1e9f;
  for (int it = 0; it < 6; ++it) {
    CK(cudaMemsetAsync(flush, it, flush_bytes));
    CK(cudaEventRecord(e0));
    probe<STAGES><<<grid, 64, smem>>>(map, base, mode, steps, box_bytes, box_rows, strips);
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    if (it > 0 && ms < best) best = ms;
  }
  printf("{\"probe\": \"%s%s\", \"stages\": %d, \"box_bytes\": %d, \"grid\": %d, \"strips\": %d, \"MB\": %.1f, \"us\": %.1f, "
         "\"GBs\": %.0f}\n",
         name, one_cta_per_sm ? " [1 CTA/SM]" : "", STAGES, box_bytes, grid, strips, total_bytes / 1e6, best * 1e3, total_bytes / best / 1e6);
}

int main() {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
  encode_t enc = (encode_t)fn;
  const size_t N = 28672, K = 4096;  // gate_up int8 weights of Llama-3.1-8B: 117 MB
  uint8_t* w;
  CK(cudaMalloc(&w, N * K));
  CK(cudaMemset(w, 1, N * K));
  const size_t flush_bytes = 512ull << 20;
  uint8_t* flush;
  CK(cudaMalloc(&flush, flush_bytes));
  // mode 0: [N rows, K bytes], box 128 rows x 128 B, strip = 128-row tile, steps = K / 128
  {
    CUtensorMap m = make_map(enc, w, N, K, 128, 128, CU_TENSOR_MAP_SWIZZLE_128B);
    run<4>("tiled 128x128B (pitch 4096)", m, w, 0, K / 128, 16384, 128, N / 128, N / 128, N * K, flush, flush_bytes);
    run<6>("tiled 128x128B (pitch 4096)", m, w, 0, K / 128, 16384, 128, N / 128, N / 128, N * K, flush, flush_bytes);
    run<12>("tiled 128x128B (pitch 4096)", m, w, 0, K / 128, 16384, 128, N / 128, N / 128, N * K, flush, flush_bytes);
    run<6>("tiled 128x128B persistent", m, w, 0, K / 128, 16384, 128, N / 128, 148, N * K, flush, flush_bytes);
    run<12>("tiled 128x128B persistent", m, w, 0, K / 128, 16384, 128, N / 128, 148, N * K, flush, flush_bytes);
  }
  // mode 1: contiguous 16 KB chunks, same strips
  {
    CUtensorMap m = make_map(enc, w, N, K, 128, 128, CU_TENSOR_MAP_SWIZZLE_128B);
    run<4>("bulk 16KB contiguous", m, w, 1, K / 128, 16384, 128, N / 128, N / 128, N * K, flush, flush_bytes);
    run<6>("bulk 16KB contiguous", m, w, 1, K / 128, 16384, 128, N / 128, N / 128, N * K, flush, flush_bytes);
    run<12>("bulk 16KB contiguous", m, w, 1, K / 128, 16384, 128, N / 128, N / 128, N * K, flush, flush_bytes);
    run<12>("bulk 16KB persistent", m, w, 1, K / 128, 16384, 128, N / 128, 148, N * K, flush, flush_bytes);
    run<12>("bulk 16KB persistent x2", m, w, 1, K / 128, 16384, 128, N / 128, 296, N * K, flush, flush_bytes);
    run<4>("tiled 128x128B", m, w, 0, K / 128, 16384, 128, N / 128, N / 128, N * K, flush, flush_bytes, true);
    run<6>("tiled 128x128B", m, w, 0, K / 128, 16384, 128, N / 128, N / 128, N * K, flush, flush_bytes, true);
    run<4>("bulk 16KB", m, w, 1, K / 128, 16384, 128, N / 128, N / 128, N * K, flush, flush_bytes, true);
    run<6>("bulk 16KB", m, w, 1, K / 128, 16384, 128, N / 128, N / 128, N * K, flush, flush_bytes, true);
  }
  // mode 2: AWQ packed words [K=5120 rows, N/2 = 27648 bytes], box 64 rows x 64 B, strip = 128 channels
  {
    const size_t Ka = 5120, rowb = 27648;  // 141.6 MB
    uint8_t* a;
    CK(cudaMalloc(&a, Ka * rowb));
    CK(cudaMemset(a, 1, Ka * rowb));
    CUtensorMap m = make_map(enc, a, Ka, rowb, 64, 64, CU_TENSOR_MAP_SWIZZLE_64B);
    run<8>("awq 64x64B (pitch 27648)", m, a, 2, Ka / 64, 4096, 64, rowb / 64, rowb / 64, Ka * rowb, flush, flush_bytes);
    run<16>("awq 64x64B (pitch 27648)", m, a, 2, Ka / 64, 4096, 64, rowb / 64, rowb / 64, Ka * rowb, flush, flush_bytes);
    run<16>("awq bulk 4KB contiguous", m, a, 1, Ka / 64, 4096, 64, rowb / 64, rowb / 64, Ka * rowb, flush, flush_bytes);
    run<32>("awq bulk 4KB contiguous", m, a, 1, Ka / 64, 4096, 64, rowb / 64, rowb / 64, Ka * rowb, flush, flush_bytes);
    run<8>("awq 64x64B", m, a, 2, Ka / 64, 4096, 64, rowb / 64, rowb / 64, Ka * rowb, flush, flush_bytes, true);
    run<16>("awq 64x64B", m, a, 2, Ka / 64, 4096, 64, rowb / 64, rowb / 64, Ka * rowb, flush, flush_bytes, true);
    run<24>("awq 64x64B", m, a, 2, Ka / 64, 4096, 64, rowb / 64, rowb / 64, Ka * rowb, flush, flush_bytes, true);
    run<8>("awq bulk 4KB", m, a, 1, Ka / 64, 4096, 64, rowb / 64, rowb / 64, Ka * rowb, flush, flush_bytes, true);
    run<16>("awq bulk 4KB", m, a, 1, Ka / 64, 4096, 64, rowb / 64, rowb / 64, Ka * rowb, flush, flush_bytes, true);
    run<24>("awq bulk 4KB", m, a, 1, Ka / 64, 4096, 64, rowb / 64, rowb / 64, Ka * rowb, flush, flush_bytes, true);
  }
  return 0;
}
