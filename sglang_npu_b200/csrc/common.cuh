// Shared device/host helpers for the b200 hot-path kernels (sm_100a only).
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>

#include "../../include/b200_hotpath.h"

#define B200_API extern "C" __attribute__((visibility("default")))

namespace b200 {

// ---- host-side error plumbing ----------------------------------------------------------
void set_error(const char* fmt, ...);
int fail(b200_status code, const char* fmt, ...);

#define B200_REQUIRE(cond, ...)                                   \
  do {                                                            \
    if (!(cond)) return ::b200::fail(B200_ERR_INVALID_ARG, __VA_ARGS__); \
  } while (0)

#define B200_CUDA_OK(expr)                                                          \
  do {                                                                              \
    cudaError_t _e = (expr);                                                        \
    if (_e != cudaSuccess)                                                          \
      return ::b200::fail(B200_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,            \
                          cudaGetErrorString(_e), __FILE__, __LINE__);              \
  } while (0)

#define B200_LAUNCH_OK()                                                            \
  do {                                                                              \
    cudaError_t _e = cudaPeekAtLastError();                                         \
    if (_e != cudaSuccess)                                                          \
      return ::b200::fail(B200_ERR_CUDA, "kernel launch failed: %s (%s:%d)",        \
                          cudaGetErrorString(_e), __FILE__, __LINE__);              \
  } while (0)

int sm_count();

// Optional per-CTA phase trace (b200_debug_set_trace): 16 x u64 per CTA — slot 0 / 15 = %globaltimer at entry / exit,
// slots 1..14 = %clock64 at kernel-specific points.  Null in normal operation (one predictable branch per point).
extern unsigned long long* g_trace;
extern size_t g_trace_slots;

// ---- programmatic dependent launch (PDL) ---------------------------------------------------
// Every kernel of the decode chain is launched with cudaLaunchAttributeProgrammaticStreamSerialization, so
// its CTAs may become resident while the previous kernel in the stream is still draining.  The contract
// inside each kernel:
//   * pdl_trigger() early: dependents may be scheduled once every CTA of this grid has started;
//   * pdl_wait() before the first access to ANYTHING a previous kernel may have written (or may still
//     be reading, for buffers this kernel overwrites) — it returns when all earlier grids have completed
//     and their writes are visible.  Only launch-invariant data (weights, cos/sin tables, the page table
//     of the step) may be touched before it: that is what lets a GEMM stream its first weight tiles, and
//     set up barriers / TMEM, under the tail of its predecessor.
// Every thread calls pdl_wait() on every path (also early-exit ones): a grid's completion must imply
// the completion of its predecessors, or the NEXT grid's pdl_wait() would not cover them.
// Both instructions are no-ops when the kernel was launched without the attribute (B200_PDL=0).
bool pdl_enabled();

template <typename... KArgs, typename... Args>
static inline int launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                           unsigned cluster_z, bool pdl, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int n = 0;
  if (cluster_z > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = 1;
    attr[n].val.clusterDim.y = 1;
    attr[n].val.clusterDim.z = cluster_z;
    ++n;
  }
  if (pdl && pdl_enabled()) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  B200_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, KArgs(args)...));
  return B200_OK;
}

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

// ---- dtype traits ----------------------------------------------------------------------
template <typename T>
struct DT;
template <>
struct DT<__nv_bfloat16> {
  using T2 = __nv_bfloat162;
  static __device__ __forceinline__ float to_f(__nv_bfloat16 v) { return __bfloat162float(v); }
  static __device__ __forceinline__ __nv_bfloat16 from_f(float v) { return __float2bfloat16_rn(v); }
  static __device__ __forceinline__ float2 to_f2(T2 v) { return __bfloat1622float2(v); }
  static __device__ __forceinline__ T2 from_f2(float a, float b) { return __floats2bfloat162_rn(a, b); }
};
template <>
struct DT<__half> {
  using T2 = __half2;
  static __device__ __forceinline__ float to_f(__half v) { return __half2float(v); }
  static __device__ __forceinline__ __half from_f(float v) { return __float2half_rn(v); }
  static __device__ __forceinline__ float2 to_f2(T2 v) { return __half22float2(v); }
  static __device__ __forceinline__ T2 from_f2(float a, float b) { return __floats2half2_rn(a, b); }
};

template <typename T>
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  typename DT<T>::T2 v = DT<T>::from_f2(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
template <typename T>
__device__ __forceinline__ float2 unpack2(uint32_t u) {
  typename DT<T>::T2 v = *reinterpret_cast<typename DT<T>::T2*>(&u);
  return DT<T>::to_f2(v);
}

// ---- small PTX wrappers ----------------------------------------------------------------
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;\n" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;\n" ::: "memory"); }

__device__ __forceinline__ void trace_mark(unsigned long long* slot16, int i) {
  if (slot16 != nullptr) {
    unsigned long long t;
    if (i == 0 || i == 15) asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t));
    else asm volatile("mov.u64 %0, %%clock64;\n" : "=l"(t));
    slot16[i] = t;
  }
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void* gptr) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(smem_addr), "l"(gptr));
}
// zero-fills when !valid (src-size 0)
__device__ __forceinline__ void cp_async16_pred(uint32_t smem_addr, const void* gptr, bool valid) {
  int sz = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(smem_addr), "l"(gptr), "r"(sz));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}

__device__ __forceinline__ void ldmatrix_x4(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3,
                                            uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t& r0, uint32_t& r1, uint32_t& r2,
                                                  uint32_t& r3, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}

// D(16x8,f32) += A(16x16) * B(16x8), bf16 / f16 inputs
template <typename T>
__device__ __forceinline__ void mma_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0,
                                          uint32_t b1);
template <>
__device__ __forceinline__ void mma_16816<__nv_bfloat16>(float (&d)[4], const uint32_t (&a)[4],
                                                         uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
      "{%0,%1,%2,%3};\n"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <>
__device__ __forceinline__ void mma_16816<__half>(float (&d)[4], const uint32_t (&a)[4],
                                                  uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
      "{%0,%1,%2,%3};\n"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm volatile("ex2.approx.ftz.f32 %0, %1;\n" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ float tanh_approx(float x) {
  float y;
  asm volatile("tanh.approx.f32 %0, %1;\n" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ int4 ld_nc_v4(const void* p) {
  int4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

__device__ __forceinline__ int64_t load_index(const void* p, int64_t i, bool is64) {
  return is64 ? reinterpret_cast<const int64_t*>(p)[i]
              : static_cast<int64_t>(reinterpret_cast<const int32_t*>(p)[i]);
}

}  // namespace b200
