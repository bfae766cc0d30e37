// Entry points declared in include/b200_hotpath.h whose kernels have not landed yet.
// They fail loudly (B200_ERR_UNSUPPORTED -> NotImplementedError); nothing falls back.
#include "common.cuh"
using namespace b200;
#define PENDING(name) return fail(B200_ERR_UNSUPPORTED, name ": kernel not built in this revision")
#define API B200_API

B200_API int b200_awq_dequantize(void*, const int32_t*, const void*, const int32_t*, int64_t, int64_t, int, int, void*) {
  PENDING("awq_dequantize");
}
B200_API int b200_awq_gemm(void*, const void*, const int32_t*, const void*, const int32_t*, const void*, int64_t, int64_t,
                      int64_t, int, int, void*) { PENDING("awq_gemm"); }
