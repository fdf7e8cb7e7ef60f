// Philox4x32-10 and the Box-Muller transform, bit-compatible with cuRAND's device code
// (curand_philox4x32_x.h:149-195, curand_normal.h:70-88) as torch.randn(device='cuda') uses it.
// Shared by philox.cu and the fused batch generator in mlp_tc.cu.  Default math flags only.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200ds {

__device__ __forceinline__ uint4 philox_round(uint4 c, uint2 k) {
  const unsigned M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
  unsigned hi0 = __umulhi(M0, c.x), lo0 = M0 * c.x;
  unsigned hi1 = __umulhi(M1, c.z), lo1 = M1 * c.z;
  return make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
}

__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
  for (int r = 0; r < 9; ++r) {
    c = philox_round(c, k);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return philox_round(c, k);
}

__device__ __forceinline__ float2 box_muller(unsigned a, unsigned b) {
  const float inv = 2.3283064e-10f;                   // 2^-32
  const float inv2pi = 2.3283064e-10f * 6.2831855f;   // 2^-32 * 2pi
  float u = a * inv + (inv / 2);
  float v = b * inv2pi + (inv2pi / 2);
  float s = sqrtf(-2.0f * logf(u));
  float2 r;
  __sincosf(v, &r.x, &r.y);
  r.x *= s;
  r.y *= s;
  return r;
}

// ATen's normal_() finishes with `rand * std + mean` (std = 1, mean = 0) in fp32
// (ATen/native/cuda/DistributionTemplates.h normal_kernel): a no-op except that it turns the
// -0.0 Box-Muller produces when u rounds to exactly 1 (sqrt(-2 log 1) = -0.0) into +0.0.
__device__ __forceinline__ float normal_transform(float r) { return __fadd_rn(__fmul_rn(r, 1.0f), 0.0f); }

// launch geometry ATen would use for `numel` outputs: (grid of 256-thread blocks, Philox calls per thread)
void philox_plan(int64_t numel, int* grid, int* calls);

}  // namespace b200ds
