// Error plumbing and library info for the C-ABI.
#include <cstdlib>

#include "common.cuh"

namespace b200 {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int fail(b200_status code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return static_cast<int>(code);
}

bool pdl_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("B200_PDL");
    on = (e != nullptr && e[0] == '0') ? 0 : 1;
  }
  return on != 0;
}

int sm_count() {
  static thread_local int cached_dev = -1;
  static thread_local int cached = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached = n;
    cached_dev = dev;
  }
  return cached;
}

}  // namespace b200

B200_API int b200_abi_version(void) { return 2; }

B200_API const char* b200_last_error(void) { return b200::g_err; }

namespace b200 {
unsigned long long* g_trace = nullptr;
size_t g_trace_slots = 0;
}  // namespace b200

B200_API int b200_debug_set_trace(void* device_buffer, size_t bytes) {
  b200::g_trace = reinterpret_cast<unsigned long long*>(device_buffer);
  b200::g_trace_slots = device_buffer ? bytes / (16 * sizeof(unsigned long long)) : 0;
  return B200_OK;
}

B200_API int b200_device_info(int* sm_count, int* cc_major, int* cc_minor) {
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (sm_count) B200_CUDA_OK(cudaDeviceGetAttribute(sm_count, cudaDevAttrMultiProcessorCount, dev));
  if (cc_major) B200_CUDA_OK(cudaDeviceGetAttribute(cc_major, cudaDevAttrComputeCapabilityMajor, dev));
  if (cc_minor) B200_CUDA_OK(cudaDeviceGetAttribute(cc_minor, cudaDevAttrComputeCapabilityMinor, dev));
  return B200_OK;
}

