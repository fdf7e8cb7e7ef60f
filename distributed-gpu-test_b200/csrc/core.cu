// Library plumbing: thread-local error string, version, device queries.
#include "common.cuh"
#include <string.h>

namespace b200ds {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  set_error("CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
  return B200DS_ERR_CUDA;
}

int sm_count() {
  static int cached[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

}  // namespace b200ds

extern "C" {

const char* b200ds_last_error(void) { return b200ds::g_err; }

int b200ds_version(void) { return 100; }  // 0.1.0

int b200ds_device_sm_count(int device, int* out) {
  B200_REQUIRE(out != nullptr, "b200ds_device_sm_count: null output");
  int n = 0;
  B200_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device));
  *out = n;
  return B200DS_OK;
}

}  // extern "C"
