// Fused Adam over the flat parameter buffer: replaces torch.optim.Adam.step as called by
// model_engine.step() (train.py:85,114; ~8 multi-tensor launches over 4 tensors) and the
// gradient clear that follows.  One pass, 128-bit accesses: reads p,g,m,v and writes
// p,m,v = 28 B/param (32 B/param when zero_grad also clears g).
//
// Arithmetic follows torch/optim/adam.py _single_tensor_adam / _multi_tensor_adam:
//   m.lerp_(g, 1-b1); v.mul_(b2).addcmul_(g, g, 1-b2);
//   denom = sqrt(v)/sqrt(1-b2^t) + eps;  p.addcdiv_(m, denom, -lr/(1-b1^t))
// with the bias corrections formed in double and rounded to fp32 once, as torch does when
// it passes Python floats to fp32 kernels.
#include "adam.cuh"

namespace b200ds {

template <bool kCounted>
__global__ void __launch_bounds__(256)
adam_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
            int64_t n, AdamScalars s, double lr, double beta1, double beta2, int64_t step,
            const unsigned long long* __restrict__ step_counter, int zero_grad) {
  __shared__ float sh[2];
  const int64_t nvec = n >> 2;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  float4* p4 = reinterpret_cast<float4*>(p);
  float4* g4 = reinterpret_cast<float4*>(g);
  float4* m4 = reinterpret_cast<float4*>(m);
  float4* v4 = reinterpret_cast<float4*>(v);
  // the first vectors are requested before the block waits for thread 0's two double-precision
  // pow(): their latency and the loads' overlap instead of adding up
  int64_t i = t0;
  float4 pv, gv, mv, vv;
  if (i < nvec) { pv = p4[i]; gv = g4[i]; mv = m4[i]; vv = v4[i]; }
  if (threadIdx.x == 0) {
    double t = kCounted ? (double)(*step_counter) : (double)step;
    double bc1 = 1.0 - pow(beta1, t);
    double bc2 = 1.0 - pow(beta2, t);
    sh[0] = (float)(-(lr / bc1));
    sh[1] = (float)sqrt(bc2);
  }
  __syncthreads();
  const float neg_step_size = sh[0], bc2_sqrt = sh[1];
  while (i < nvec) {
    adam_one(pv.x, gv.x, mv.x, vv.x, s, neg_step_size, bc2_sqrt);
    adam_one(pv.y, gv.y, mv.y, vv.y, s, neg_step_size, bc2_sqrt);
    adam_one(pv.z, gv.z, mv.z, vv.z, s, neg_step_size, bc2_sqrt);
    adam_one(pv.w, gv.w, mv.w, vv.w, s, neg_step_size, bc2_sqrt);
    p4[i] = pv; m4[i] = mv; v4[i] = vv;
    if (zero_grad) g4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    i += stride;
    if (i < nvec) { pv = p4[i]; gv = g4[i]; mv = m4[i]; vv = v4[i]; }
  }
  for (int64_t i = (nvec << 2) + t0; i < n; i += stride) {
    float pv = p[i], mv = m[i], vv = v[i];
    adam_one(pv, g[i], mv, vv, s, neg_step_size, bc2_sqrt);
    p[i] = pv; m[i] = mv; v[i] = vv;
    if (zero_grad) g[i] = 0.f;
  }
}

static int launch_adam(float* p, float* g, float* m, float* v, int64_t n, double lr, double beta1,
                       double beta2, double eps, int64_t step, const uint64_t* counter, float grad_scale,
                       int zero_grad, uintptr_t stream) {
  B200_REQUIRE(n >= 0, "adam_step: negative n");
  if (n == 0) return B200DS_OK;
  B200_REQUIRE(p && g && m && v, "adam_step: null pointer");
  B200_REQUIRE(((((uintptr_t)p) | ((uintptr_t)g) | ((uintptr_t)m) | ((uintptr_t)v)) & 15) == 0,
               "adam_step: buffers must be 16-byte aligned");
  B200_REQUIRE(counter != nullptr || step >= 1, "adam_step: step must be >= 1 (t after the increment)");
  AdamScalars s;
  // torch forms 1-beta in double and hands it to an fp32 kernel
  s.w1 = (float)(1.0 - beta1);
  s.beta2 = (float)beta2;
  s.w2 = (float)(1.0 - beta2);
  s.eps = (float)eps;
  s.grad_scale = grad_scale;
  int64_t want = ((n >> 2) + 255) / 256;
  int64_t cap = (int64_t)sm_count() * 8;
  int grid = (int)(want < 1 ? 1 : (want > cap ? cap : want));
  if (counter)
    adam_kernel<true><<<grid, 256, 0, as_stream(stream)>>>(p, g, m, v, n, s, lr, beta1, beta2, 0,
                                                            reinterpret_cast<const unsigned long long*>(counter), zero_grad);
  else
    adam_kernel<false><<<grid, 256, 0, as_stream(stream)>>>(p, g, m, v, n, s, lr, beta1, beta2, step,
                                                             nullptr, zero_grad);
  B200_LAUNCH_CHECK();
  return B200DS_OK;
}

}  // namespace b200ds

using namespace b200ds;

extern "C" {

int b200ds_adam_step(float* params_dev, float* grads_dev, float* exp_avg_dev, float* exp_avg_sq_dev,
                     int64_t n, double lr, double beta1, double beta2, double eps, int64_t step,
                     float grad_scale, int zero_grad, uintptr_t stream) {
  return launch_adam(params_dev, grads_dev, exp_avg_dev, exp_avg_sq_dev, n, lr, beta1, beta2, eps, step,
                     nullptr, grad_scale, zero_grad, stream);
}

int b200ds_adam_step_counted(float* params_dev, float* grads_dev, float* exp_avg_dev,
                             float* exp_avg_sq_dev, int64_t n, double lr, double beta1, double beta2,
                             double eps, const uint64_t* step_counter_dev, float grad_scale,
                             int zero_grad, uintptr_t stream) {
  B200_REQUIRE(step_counter_dev != nullptr, "adam_step_counted: null step counter");
  return launch_adam(params_dev, grads_dev, exp_avg_dev, exp_avg_sq_dev, n, lr, beta1, beta2, eps, 0,
                     step_counter_dev, grad_scale, zero_grad, stream);
}

}  // extern "C"
