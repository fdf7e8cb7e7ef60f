// Fused whole-step kernel for the reference's own shape class (train.py:63,74,76: F=20,
// H=64, per-rank batch 4): everything train.py:104-117 does per step -- batch gather,
// forward, BCE loss, backward, Adam, the per-step loss value -- in ONE launch by ONE CTA,
// and with apply_adam the kernel keeps the 1409 parameters and their Adam state in
// registers across n_steps consecutive steps.  The reference spends ~35-40 launches, two
// blocking H2D copies and one blocking D2H per step on the same work.
//
// Thread n owns hidden unit n: W1[n,:], b1[n], W2[n] and their Adam moments live in its
// registers, so dW1/db1/dW2 and the update are thread-local; the only cross-thread
// traffic per step is the layer-2 dot product (warp shuffle + one smem hop).
// The next step's rows are fetched with cp.async while the current step computes.
#include "common.cuh"

namespace b200ds {

constexpr int kTinyMaxBatch = 32;
constexpr int kTinyMaxHidden = 256;
constexpr int kTinyMaxOrder = 8192;   // samples per launch (indices staged in smem)

__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gsrc) {
  unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(d), "l"(gsrc));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;\n" ::: "memory"); }

struct TinyArgs {
  float* params; float* grads; float* m; float* v;
  const float* x_all; const float* y_all; const int64_t* order;
  int64_t n_order, first;
  int F, H, batch, n_steps, apply_adam;
  double lr, beta1, beta2;
  long long step0;                           // Adam t before the first step of this launch
  float w1, beta2f, w2, eps;                 // (float)(1-beta1), (float)beta2, (float)(1-beta2), (float)eps
  float* losses;
};

// Adam for a register-resident parameter.  This kernel is latency-bound (one CTA, a few
// hundred dependent instructions per step), so the two divisions and the square root of the
// update use the hardware approximations (sqrt.approx / rcp, <= 2 ulp each) instead of the
// IEEE sequences adam.cu uses; the difference is ~1e-7 relative per update, two orders of
// magnitude inside the 1e-5 parity tolerance.  inv_bc2_sqrt = 1/sqrt(1 - beta2^t).
__device__ __forceinline__ void adam_reg(float& p, float g, float& m, float& v, const TinyArgs& a,
                                         float neg_step, float inv_bc2_sqrt) {
  m = fmaf(a.w1, g - m, m);
  v = fmaf(a.w2 * g, g, v * a.beta2f);
  float sq;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq) : "f"(v));
  const float denom = fmaf(sq, inv_bc2_sqrt, a.eps);
  p = fmaf(neg_step, __fdividef(m, denom), p);
}

// BT: compile-time bound on the per-step batch (row loops fully unrolled and interleaved)
template <int FP, int BT>
__global__ void __launch_bounds__(kTinyMaxHidden, 1) tiny_train_steps_kernel(TinyArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;
  const int F = a.F, H = a.H, B = a.batch, HP = blockDim.x;
  // smem carve-up
  float* xs = reinterpret_cast<float*>(smem_raw);            // [2][B][FP]
  float* ys = xs + 2 * B * FP;                                // [2][B]
  float* hs = ys + 2 * B;                                     // [B][HP]
  float* zpart = hs + B * HP;                                 // [B][nwarp]
  float* dzs = zpart + B * nwarp;                             // [B]
  float* terms = dzs + B;                                     // [B]
  float* sc = terms + B;                                      // [n_steps][2] adam scalars (-lr/bc1, 1/sqrt(bc2)) per step
  int* ord = reinterpret_cast<int*>(sc + 2 * a.n_steps);      // [n_steps*B]

  const Layout L{F, H};
  const bool own = tid < H;
  float w[FP], mw[FP], vw[FP];
  float b1 = 0.f, w2 = 0.f, b2 = 0.f, mb1 = 0.f, vb1 = 0.f, mw2 = 0.f, vw2 = 0.f, mb2 = 0.f, vb2 = 0.f;
#pragma unroll
  for (int f = 0; f < FP; ++f) {
    bool ok = own && f < F;
    w[f] = ok ? a.params[L.off_w1() + (int64_t)tid * F + f] : 0.f;
    mw[f] = (ok && a.apply_adam) ? a.m[L.off_w1() + (int64_t)tid * F + f] : 0.f;
    vw[f] = (ok && a.apply_adam) ? a.v[L.off_w1() + (int64_t)tid * F + f] : 0.f;
  }
  if (own) {
    b1 = a.params[L.off_b1() + tid];
    w2 = a.params[L.off_w2() + tid];
    if (a.apply_adam) {
      mb1 = a.m[L.off_b1() + tid]; vb1 = a.v[L.off_b1() + tid];
      mw2 = a.m[L.off_w2() + tid]; vw2 = a.v[L.off_w2() + tid];
    }
  }
  b2 = a.params[L.off_b2()];
  if (a.apply_adam) { mb2 = a.m[L.off_b2()]; vb2 = a.v[L.off_b2()]; }   // every thread tracks b2

  const int64_t avail = a.n_order - a.first;
  const int total = (int)min((int64_t)a.n_steps * B, avail);
  for (int i = tid; i < total; i += blockDim.x) ord[i] = (int)a.order[a.first + i];
  for (int i = tid; i < 2 * B * FP; i += blockDim.x) xs[i] = 0.f;   // zero the K padding once
  if (a.apply_adam) {
    // bias corrections of every step of this launch, formed in double exactly as torch does
    // (1 - beta ** t), in parallel and off the per-step critical path
    for (int i = tid; i < a.n_steps; i += blockDim.x) {
      const double t = (double)(a.step0 + i + 1);
      sc[2 * i + 0] = (float)(-(a.lr / (1.0 - pow(a.beta1, t))));
      sc[2 * i + 1] = (float)(1.0 / sqrt(1.0 - pow(a.beta2, t)));
    }
  }
  __syncthreads();

  auto prefetch = [&](int s) {   // rows of step s -> buffer s&1
    const int pos = s * B;
    const int nb = min(B, total - pos);
    float* xb = xs + (s & 1) * B * FP;
    float* yb = ys + (s & 1) * B;
    for (int i = tid; i < nb * F; i += blockDim.x) {
      int r = i / F, f = i - r * F;
      cp_async4(xb + r * FP + f, a.x_all + (int64_t)ord[pos + r] * F + f);
    }
    if (tid < nb) cp_async4(yb + tid, a.y_all + ord[pos + tid]);
    cp_async_commit();
  };

  if (total > 0) prefetch(0);
  for (int s = 0; s < a.n_steps; ++s) {
    const int pos = s * B;
    const int nb = min(B, total - pos);
    if (nb <= 0) break;
    cp_async_wait_all();
    __syncthreads();                       // S1: step s rows landed; everyone is done with step s-1
    if (s + 1 < a.n_steps && pos + B < total) prefetch(s + 1);
    const float* xb = xs + (s & 1) * B * FP;
    const float* yb = ys + (s & 1) * B;

    // ---- forward: z1 = b1 + x . W1[n,:], h = relu(z1), partial of h . w2
    {
      float hrow[BT], pz[BT];
#pragma unroll
      for (int r = 0; r < BT; ++r) {
        float z1 = b1;
        if (r < nb) {
          const float4* xr = reinterpret_cast<const float4*>(xb + r * FP);
#pragma unroll
          for (int q = 0; q < FP / 4; ++q) {
            float4 xv = xr[q];
            z1 = fmaf(xv.x, w[4 * q + 0], z1);
            z1 = fmaf(xv.y, w[4 * q + 1], z1);
            z1 = fmaf(xv.z, w[4 * q + 2], z1);
            z1 = fmaf(xv.w, w[4 * q + 3], z1);
          }
        }
        hrow[r] = (own && r < nb) ? fmaxf(z1, 0.f) : 0.f;
        pz[r] = hrow[r] * w2;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int r = 0; r < BT; ++r) pz[r] += __shfl_xor_sync(0xffffffffu, pz[r], o);
      }
#pragma unroll
      for (int r = 0; r < BT; ++r) {
        if (r < nb) {
          hs[r * HP + tid] = hrow[r];
          if (lane == 0) zpart[r * nwarp + warp] = pz[r];
        }
      }
    }
    __syncthreads();                       // S2
    {   // loss term and its derivative of row r are formed by two different warps side by side
      const bool second = nwarp > 1;
      const int r = second ? (tid & 31) : tid;
      const bool do_term = tid < nb, do_grad = second ? (warp == 1 && r < nb) : (tid < nb);
      if (do_term || do_grad) {
        float z = 0.f;
        for (int q = 0; q < nwarp; ++q) z += zpart[r * nwarp + q];
        z += b2;
        const float y = yb[r];
        if (do_term) terms[r] = bce_term(z, y);
        if (do_grad) dzs[r] = bce_grad(z, y, (float)nb);
      }
    }
    __syncthreads();                       // S3
    if (tid == 0) {
      float t = 0.f;
      for (int r = 0; r < nb; ++r) t += terms[r];
      a.losses[s] = t / (float)nb;
    }

    // ---- backward: thread-local gradients of the unit this thread owns
    float gw[FP];
#pragma unroll
    for (int f = 0; f < FP; ++f) gw[f] = 0.f;
    float gb1 = 0.f, gw2 = 0.f, gb2 = 0.f;
#pragma unroll
    for (int r = 0; r < BT; ++r) {
      if (r >= nb) break;
      float d = dzs[r];
      float h = hs[r * HP + tid];
      float gz = h > 0.f ? d * w2 : 0.f;
      gb2 += d;
      gb1 += gz;
      gw2 = fmaf(d, h, gw2);
      const float4* xr = reinterpret_cast<const float4*>(xb + r * FP);
#pragma unroll
      for (int q = 0; q < FP / 4; ++q) {
        float4 xv = xr[q];
        gw[4 * q + 0] = fmaf(gz, xv.x, gw[4 * q + 0]);
        gw[4 * q + 1] = fmaf(gz, xv.y, gw[4 * q + 1]);
        gw[4 * q + 2] = fmaf(gz, xv.z, gw[4 * q + 2]);
        gw[4 * q + 3] = fmaf(gz, xv.w, gw[4 * q + 3]);
      }
    }

    if (a.apply_adam) {
      const float neg_step = sc[2 * s + 0], inv_bc2 = sc[2 * s + 1];
      // Unpredicated and staged: padding columns and threads beyond H hold zeros everywhere
      // (g = m = v = 0 leaves p untouched), so every thread runs the same straight-line code and
      // the FP + 3 independent updates overlap their sqrt / reciprocal latencies.
      float sq[FP], qd[FP];
#pragma unroll
      for (int f = 0; f < FP; ++f) {
        mw[f] = fmaf(a.w1, gw[f] - mw[f], mw[f]);
        vw[f] = fmaf(a.w2 * gw[f], gw[f], vw[f] * a.beta2f);
      }
#pragma unroll
      for (int f = 0; f < FP; ++f) asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq[f]) : "f"(vw[f]));
#pragma unroll
      for (int f = 0; f < FP; ++f) qd[f] = __fdividef(mw[f], fmaf(sq[f], inv_bc2, a.eps));
#pragma unroll
      for (int f = 0; f < FP; ++f) w[f] = fmaf(neg_step, qd[f], w[f]);
      adam_reg(b1, gb1, mb1, vb1, a, neg_step, inv_bc2);
      adam_reg(w2, gw2, mw2, vw2, a, neg_step, inv_bc2);
      // every thread tracks b2 identically (it is needed by the threads that form z)
      adam_reg(b2, gb2, mb2, vb2, a, neg_step, inv_bc2);
    } else {
      if (own) {
#pragma unroll
        for (int f = 0; f < FP; ++f)
          if (f < F) a.grads[L.off_w1() + (int64_t)tid * F + f] = gw[f];
        a.grads[L.off_b1() + tid] = gb1;
        a.grads[L.off_w2() + tid] = gw2;
      }
      if (tid == 0) a.grads[L.off_b2()] = gb2;
    }
  }

  if (a.apply_adam) {
    if (own) {
#pragma unroll
      for (int f = 0; f < FP; ++f)
        if (f < F) {
          a.params[L.off_w1() + (int64_t)tid * F + f] = w[f];
          a.m[L.off_w1() + (int64_t)tid * F + f] = mw[f];
          a.v[L.off_w1() + (int64_t)tid * F + f] = vw[f];
        }
      a.params[L.off_b1() + tid] = b1; a.m[L.off_b1() + tid] = mb1; a.v[L.off_b1() + tid] = vb1;
      a.params[L.off_w2() + tid] = w2; a.m[L.off_w2() + tid] = mw2; a.v[L.off_w2() + tid] = vw2;
    }
    if (tid == 0) { a.params[L.off_b2()] = b2; a.m[L.off_b2()] = mb2; a.v[L.off_b2()] = vb2; }
  }
}

template <int FP, int BT>
static int launch_tiny_bt(const TinyArgs& a, cudaStream_t st) {
  int threads = (a.H + 31) / 32 * 32;
  int nwarp = threads / 32;
  int total = (int)((int64_t)a.n_steps * a.batch);
  size_t smem = sizeof(float) * ((size_t)2 * a.batch * FP + 2 * a.batch + (size_t)a.batch * threads +
                                 (size_t)a.batch * nwarp + 2 * a.batch + 2 * (size_t)a.n_steps) + sizeof(int) * (size_t)total;
  smem = (smem + 15) / 16 * 16;
  if (smem > 48 * 1024)
    B200_CUDA(cudaFuncSetAttribute(tiny_train_steps_kernel<FP, BT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  tiny_train_steps_kernel<FP, BT><<<1, threads, smem, st>>>(a);
  B200_LAUNCH_CHECK();
  return B200DS_OK;
}

template <int FP>
static int launch_tiny(const TinyArgs& a, cudaStream_t st) {
  if (a.batch <= 4) return launch_tiny_bt<FP, 4>(a, st);
  if (a.batch <= 8) return launch_tiny_bt<FP, 8>(a, st);
  return launch_tiny_bt<FP, kTinyMaxBatch>(a, st);
}

}  // namespace b200ds

using namespace b200ds;

extern "C" {

int b200ds_tiny_train_steps_f32(float* params_dev, float* grads_dev, float* exp_avg_dev,
                                float* exp_avg_sq_dev, int n_features, int hidden,
                                const float* x_all_dev, const float* y_all_dev,
                                const int64_t* order_dev, int64_t n_order, int64_t first,
                                int batch, int n_steps, int apply_adam, double lr, double beta1,
                                double beta2, double eps, int64_t step0, float* losses_dev,
                                uintptr_t stream) {
  B200_REQUIRE(n_features >= 1 && n_features <= 32, "tiny_train_steps: F must be in [1,32] (got %d)", n_features);
  B200_REQUIRE(hidden >= 1 && hidden <= kTinyMaxHidden, "tiny_train_steps: H must be in [1,%d] (got %d)", kTinyMaxHidden, hidden);
  B200_REQUIRE(batch >= 1 && batch <= kTinyMaxBatch, "tiny_train_steps: batch must be in [1,%d] (got %d)", kTinyMaxBatch, batch);
  B200_REQUIRE(n_steps >= 0, "tiny_train_steps: negative n_steps");
  if (n_steps == 0) return B200DS_OK;
  B200_REQUIRE((int64_t)n_steps * batch <= kTinyMaxOrder, "tiny_train_steps: n_steps*batch must be <= %d per launch", kTinyMaxOrder);
  B200_REQUIRE(first >= 0 && first < n_order, "tiny_train_steps: first=%lld outside order[0,%lld)", (long long)first, (long long)n_order);
  B200_REQUIRE(params_dev && x_all_dev && y_all_dev && order_dev && losses_dev, "tiny_train_steps: null pointer");
  if (apply_adam) {
    B200_REQUIRE(exp_avg_dev && exp_avg_sq_dev, "tiny_train_steps: Adam state required with apply_adam");
    B200_REQUIRE(step0 >= 0, "tiny_train_steps: negative step0");
  } else {
    B200_REQUIRE(n_steps == 1, "tiny_train_steps: without apply_adam exactly one step per launch");
    B200_REQUIRE(grads_dev != nullptr, "tiny_train_steps: grads buffer required without apply_adam");
  }
  TinyArgs a;
  a.params = params_dev; a.grads = grads_dev; a.m = exp_avg_dev; a.v = exp_avg_sq_dev;
  a.x_all = x_all_dev; a.y_all = y_all_dev; a.order = order_dev;
  a.n_order = n_order; a.first = first;
  a.F = n_features; a.H = hidden; a.batch = batch; a.n_steps = n_steps; a.apply_adam = apply_adam;
  a.lr = lr; a.beta1 = beta1; a.beta2 = beta2;
  a.step0 = step0;
  a.w1 = (float)(1.0 - beta1); a.beta2f = (float)beta2; a.w2 = (float)(1.0 - beta2); a.eps = (float)eps;
  a.losses = losses_dev;
  cudaStream_t st = as_stream(stream);
  int fp = (n_features + 3) / 4 * 4;
  switch (fp) {
    case 4: return launch_tiny<4>(a, st);
    case 8: return launch_tiny<8>(a, st);
    case 12: return launch_tiny<12>(a, st);
    case 16: return launch_tiny<16>(a, st);
    case 20: return launch_tiny<20>(a, st);
    case 24: return launch_tiny<24>(a, st);
    case 28: return launch_tiny<28>(a, st);
    default: return launch_tiny<32>(a, st);
  }
}

}  // extern "C"
