// Fused BCE-with-logits loss + d(loss)/d(logits): replaces nn.BCEWithLogitsLoss()
// (train.py:96,112; ~5 ATen launches) and the first autograd node of
// model_engine.backward (train.py:113; 5 more).  One pass over z and y: 8 B read + 4 B
// written per sample, warp-shuffle reductions, one atomic ticket per CTA so the mean is
// finished by the last CTA in a fixed order (deterministic, single launch).
#include "common.cuh"

namespace b200ds {

constexpr int kLossThreads = 256;
constexpr int kLossMaxBlocks = 1184;   // 8 CTAs x 148 SMs

struct LossScratch {          // caller-provided, zeroed once; the kernel leaves ticket == 0
  unsigned int ticket;
  unsigned int pad[3];
  float partial[kLossMaxBlocks];
};

__global__ void __launch_bounds__(kLossThreads)
bce_loss_grad_kernel(const float* __restrict__ z, const float* __restrict__ y, int64_t batch,
                     float* __restrict__ loss, float* __restrict__ dz,
                     unsigned long long* __restrict__ step_counter, LossScratch* __restrict__ scr) {
  __shared__ float warp_part[kLossThreads / 32];
  __shared__ bool is_last;
  const float fb = (float)batch;
  float acc = 0.f;
  // 128-bit main body, scalar tail
  const int64_t nvec = batch >> 2;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool vec_ok = ((((uintptr_t)z) | ((uintptr_t)y) | ((uintptr_t)dz)) & 15) == 0;
  if (vec_ok) {
    for (int64_t i = t0; i < nvec; i += stride) {
      float4 zv = __ldg(reinterpret_cast<const float4*>(z) + i);
      float4 yv = __ldg(reinterpret_cast<const float4*>(y) + i);
      acc += bce_term(zv.x, yv.x);
      acc += bce_term(zv.y, yv.y);
      acc += bce_term(zv.z, yv.z);
      acc += bce_term(zv.w, yv.w);
      if (dz) {
        float4 g = make_float4(bce_grad(zv.x, yv.x, fb), bce_grad(zv.y, yv.y, fb),
                               bce_grad(zv.z, yv.z, fb), bce_grad(zv.w, yv.w, fb));
        reinterpret_cast<float4*>(dz)[i] = g;
      }
    }
    for (int64_t i = (nvec << 2) + t0; i < batch; i += stride) {
      float zi = z[i], yi = y[i];
      acc += bce_term(zi, yi);
      if (dz) dz[i] = bce_grad(zi, yi, fb);
    }
  } else {
    for (int64_t i = t0; i < batch; i += stride) {
      float zi = z[i], yi = y[i];
      acc += bce_term(zi, yi);
      if (dz) dz[i] = bce_grad(zi, yi, fb);
    }
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) warp_part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < kLossThreads / 32 ? warp_part[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) {
      if (gridDim.x == 1) {
        *loss = v / fb;
        if (step_counter) *step_counter += 1ull;
        is_last = false;
      } else {
        scr->partial[blockIdx.x] = v;
        __threadfence();
        unsigned t = atomicAdd(&scr->ticket, 1u);
        is_last = (t == gridDim.x - 1);
      }
    }
  }
  __syncthreads();
  if (is_last) {   // the CTA that arrived last sums the partials in block order, in double
    __threadfence();
    double s = 0.0;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x) s += (double)__ldcg(&scr->partial[i]);
    __shared__ double dpart[kLossThreads / 32];
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) dpart[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
      double tot = 0.0;
      for (int w = 0; w < kLossThreads / 32; ++w) tot += dpart[w];
      *loss = (float)(tot / (double)batch);
      if (step_counter) *step_counter += 1ull;
      scr->ticket = 0u;
    }
  }
}

}  // namespace b200ds

using namespace b200ds;

extern "C" {

size_t b200ds_bce_scratch_bytes(int64_t) { return sizeof(LossScratch); }

int b200ds_bce_loss_grad(const float* logits_dev, const float* y_dev, int64_t batch,
                         float* loss_dev, float* dlogits_dev, uint64_t* step_counter_dev,
                         void* scratch_dev, size_t scratch_bytes, uintptr_t stream) {
  B200_REQUIRE(batch > 0, "bce_loss_grad: batch must be positive (got %lld)", (long long)batch);
  B200_REQUIRE(logits_dev && y_dev && loss_dev, "bce_loss_grad: null pointer");
  int64_t want = (batch / 4 + kLossThreads - 1) / kLossThreads;
  int grid = (int)(want < 1 ? 1 : (want > kLossMaxBlocks ? kLossMaxBlocks : want));
  if (grid > 1)
    B200_REQUIRE(scratch_dev && scratch_bytes >= sizeof(LossScratch),
                 "bce_loss_grad: scratch of %zu bytes required (zeroed once)", sizeof(LossScratch));
  bce_loss_grad_kernel<<<grid, kLossThreads, 0, as_stream(stream)>>>(
      logits_dev, y_dev, batch, loss_dev, dlogits_dev,
      reinterpret_cast<unsigned long long*>(step_counter_dev), reinterpret_cast<LossScratch*>(scratch_dev));
  B200_LAUNCH_CHECK();
  return B200DS_OK;
}

}  // extern "C"
