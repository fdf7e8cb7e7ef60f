// The Adam arithmetic of adam.cu, shared with the fused all-reduce + Adam kernel (p2p.cu).
#pragma once
#include "common.cuh"

namespace b200ds {

struct AdamScalars {
  float w1;         // 1 - beta1        (lerp weight)
  float beta2;
  float w2;         // 1 - beta2
  float eps;
  float grad_scale; // 1/world (stage-0 averaging folded in), 1 otherwise
};

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, const AdamScalars& s,
                                         float neg_step_size, float bc2_sqrt) {
  g *= s.grad_scale;
  m = fmaf(s.w1, g - m, m);                 // lerp, weight < 0.5 branch
  v = v * s.beta2;
  v = fmaf(s.w2 * g, g, v);                 // addcmul: v + (value*g)*g
  float denom = sqrtf(v) / bc2_sqrt + s.eps;
  p = fmaf(neg_step_size, m / denom, p);    // addcdiv: p + value*(m/denom)
}

// torch forms 1-beta in double and hands it to an fp32 kernel
inline AdamScalars make_adam_scalars(double beta1, double beta2, double eps, float grad_scale) {
  AdamScalars s;
  s.w1 = (float)(1.0 - beta1);
  s.beta2 = (float)beta2;
  s.w2 = (float)(1.0 - beta2);
  s.eps = (float)eps;
  s.grad_scale = grad_scale;
  return s;
}

}  // namespace b200ds
