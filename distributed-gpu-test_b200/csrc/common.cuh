// Shared helpers for the b200ds kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/b200ds.h"

namespace b200ds {

// ---- error plumbing: nothing throws across the C boundary -------------------------
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define B200_CUDA(expr)                                                            \
  do {                                                                             \
    cudaError_t _e = (expr);                                                       \
    if (_e != cudaSuccess) return ::b200ds::cuda_fail(_e, #expr, __FILE__, __LINE__); \
  } while (0)

#define B200_REQUIRE(cond, ...)                                                    \
  do {                                                                             \
    if (!(cond)) { ::b200ds::set_error(__VA_ARGS__); return B200DS_ERR_ARG; }      \
  } while (0)

#define B200_LAUNCH_CHECK() B200_CUDA(cudaGetLastError())

static inline cudaStream_t as_stream(uintptr_t s) { return reinterpret_cast<cudaStream_t>(s); }

// SM count of the current device (cached per device).
int sm_count();

// ---- flat parameter layout: [W1 (H*F) | b1 (H) | W2 (H) | b2 (1)] -----------------
struct Layout {
  int F, H;
  __host__ __device__ int64_t off_w1() const { return 0; }
  __host__ __device__ int64_t off_b1() const { return (int64_t)H * F; }
  __host__ __device__ int64_t off_w2() const { return (int64_t)H * F + H; }
  __host__ __device__ int64_t off_b2() const { return (int64_t)H * F + 2 * (int64_t)H; }
  __host__ __device__ int64_t count() const { return (int64_t)H * F + 2 * (int64_t)H + 1; }
};

#ifdef __CUDACC__
// ---- warp / block reductions -------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---- the loss arithmetic, one definition used by every kernel ----------------------
// l = (1-y) z - logsigmoid(z), logsigmoid(z) = min(z,0) - log1p(exp(-|z|))
// (ATen binary_cross_entropy_with_logits + log_sigmoid_forward)
__device__ __forceinline__ float bce_term(float z, float y) {
  float ls = fminf(z, 0.f) - log1pf(expf(-fabsf(z)));
  return (1.f - y) * z - ls;
}
// dz = (sigmoid(z) - y) / B   (ATen sigmoid, sub_, mul_(grad=1), div_(numel))
__device__ __forceinline__ float bce_grad(float z, float y, float batch) {
  float s = 1.f / (1.f + expf(-z));
  return (s - y) / batch;
}
#endif

}  // namespace b200ds
